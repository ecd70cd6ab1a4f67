"""Fused o_proj GEMM + all-reduce across a TP group: parity and timing.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        benchmarks/multi/oproj_ar.py [--json out.json] [--iters 50]

Parity: every rank's output equals oproj_allreduce_ref (partials rounded to the model dtype, summed in fp32)
within one ulp, is bit-identical on all ranks, for the multicast (NVSwitch multimem) and the peer-load path.
Timing (CUDA events on the launching stream, max over ranks): the fused kernel vs the reference's data flow --
cuBLAS GEMM (torch.matmul) followed by an NCCL all-reduce -- on the Llama-70B o_proj shard shapes of cfg4."""
import argparse
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

REPO = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(REPO))
from sglang_b200.comm import OProjAllReduce, oproj_allreduce_ref  # noqa: E402


def timed(fn, iters, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters * 1e3], device="cuda")       # us
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def timed_graph(fn, reps=4, iters=10):
    """Device time of one call: `reps` calls captured in a CUDA graph and replayed (what a serving engine does for
    decode), which takes the Python / ctypes launch cost out of the number."""
    fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(reps):
                fn()
    torch.cuda.current_stream().wait_stream(side)
    return timed(g.replay, iters, warmup=2) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default="")
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--graph", type=int, default=1, help="time through CUDA-graph replays (1) or eager launches (0)")
    ap.add_argument("--parity-only", action="store_true")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    dt = torch.bfloat16
    results = {"world": world, "parity": [], "timing": []}
    ok = True

    # ---------------------------------------------------------------- parity
    for (T, N, K_full) in [(1, 64, 128 * world), (64, 8192, 8192), (37, 4096, 4096), (128, 1024, 512 * world),
                           (200, 2048, 1024 * world), (777, 8192, 8192)]:
        K = K_full // world
        g = torch.Generator().manual_seed(1000 + T + N)             # same seed everywhere: every rank knows all shards
        xs = [(torch.randn(T, K, generator=g) * 0.5).to(dt) for _ in range(world)]
        ws = [(torch.randn(N, K, generator=g) * 0.5).to(dt) for _ in range(world)]
        want, mag = oproj_allreduce_ref(xs, ws, with_scale=True)
        want = want.float()
        x, w = xs[rank].cuda(), ws[rank].cuda()
        for mc, mode in ([(m, o) for m in (True, False) for o in (1, 2)] if world > 1 else [(False, 0)]):
            comm = OProjAllReduce(dist.group.WORLD, max(T, 128), N, dt, "cuda", use_multicast=mc, mode=mode)
            for rep in range(4):                                    # repeated calls reuse buffers and flags
                got = comm.forward(x, w).clone()
            torch.cuda.synchronize()
            replayed = False
            if T <= 128:                                            # the device-side launch counter under graph replay
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    gr = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gr, stream=side):
                        g_out = comm.forward(x, w).clone()
                        g_out2 = comm.forward(x, w).clone()
                torch.cuda.current_stream().wait_stream(side)
                for rep in range(3):
                    gr.replay()
                torch.cuda.synchronize()
                replayed = bool(torch.equal(g_out, got) and torch.equal(g_out2, got))
                got = g_out2
                del gr
            err = (got.float().cpu() - want).abs()
            # bf16 ulp is 2^-7 relative; each partial and the final sum may round one ulp either way (fp32 vs
            # float64 accumulation), so allow two ulps of the partials' magnitude and report how many are exact
            tol = 2.0 ** -7 * mag.clamp_min(1.0) * 2
            good = bool((err <= tol).all())
            exact = float((err == 0).float().mean())
            gathered = [torch.empty_like(got) for _ in range(world)]
            dist.all_gather(gathered, got)
            same = all(torch.equal(gathered[0], t) for t in gathered)
            ok = ok and good and same and (replayed or T > 128)
            if rank == 0:
                rec = dict(T=T, N=N, K_shard=K, multicast=bool(mc and comm.multicast),
                           mode={0: "auto", 1: "one-shot", 2: "two-shot"}[mode], max_err=float(err.max()),
                           exact_fraction=round(exact, 5),
                           within_ulp=good, identical_on_all_ranks=same, graph_replay=replayed)
                results["parity"].append(rec)
                print(json.dumps(rec), flush=True)
            del comm

    # ---------------------------------------------------------------- timing: Llama-70B o_proj shard (hidden 8192)
    for T in (() if a.parity_only else (8, 32, 64, 128, 512, 2048)):
        N, K = 8192, 8192 // world
        x = (torch.randn(T, K, device="cuda") * 0.5).to(dt)
        # the weights rotate through copies larger than L2, so every GEMM streams them from HBM like a layer stack
        n_copies = max(2, (300 << 20) // (N * K * 2) + 1)
        ws = [(torch.randn(N, K, device="cuda") * 0.5).to(dt) for _ in range(n_copies)]
        out = torch.empty(T, N, dtype=dt, device="cuda")
        rec = dict(T=T, N=N, K_shard=K, weight_copies=n_copies)

        def baseline():
            for wi in ws:
                torch.matmul(x, wi.t(), out=out)
                if world > 1:
                    dist.all_reduce(out)

        def gemm_only():
            for wi in ws:
                torch.matmul(x, wi.t(), out=out)

        tm = timed_graph if a.graph else (lambda fn: timed(fn, a.iters))
        rec["timing"] = "cuda-graph replay" if a.graph else "eager"
        rec["cublas_gemm_us"] = round(tm(gemm_only) / n_copies, 2)
        rec["cublas_plus_nccl_us"] = round(tm(baseline) / n_copies, 2)
        for mc, mode in ([(m, o) for m in (True, False) for o in (1, 2)] if world > 1 else [(False, 0)]):
            comm = OProjAllReduce(dist.group.WORLD, max(T, 128), N, dt, "cuda", use_multicast=mc, mode=mode)
            key = "fused_" + ("multimem" if (mc and comm.multicast) else "peer") + {0: "", 1: "_1shot", 2: "_2shot"}[mode] + "_us"
            rec[key] = round(tm(lambda: comm.forward(x, ws[0]) if False else [comm.forward(x, wi) for wi in ws]), 2) / len(ws)
            rec[key] = round(rec[key], 2)
            del comm
        bytes_w = N * K * 2 + T * K * 2 + T * N * 2
        best = min(v for k, v in rec.items() if k.startswith("fused"))
        rec["fused_gbs_weight_stream"] = round(bytes_w / best / 1e3, 1)
        rec["speedup_vs_cublas_nccl"] = round(rec["cublas_plus_nccl_us"] / best, 2)
        if rank == 0:
            results["timing"].append(rec)
            print(json.dumps(rec), flush=True)

    if rank == 0:
        results["ok"] = ok
        print("PARITY", "OK" if ok else "FAILED", flush=True)
        if a.json:
            Path(a.json).write_text(json.dumps(results, indent=1))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
