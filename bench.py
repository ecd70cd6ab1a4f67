#!/usr/bin/env python
"""bench.py -- RadixAttention hot-path benchmark on B200 (contract: DESIGN.md section 5).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

N = 1  (BASELINE.json configs[1], "cfg2"): Llama-3.1-8B shapes, bf16, TP1, paged decode, batch 64 at 4K context,
       page_size 1, random slot permutation.  One step = the attention path of one decode step of all 32 layers:
       1 planner launch + 32 x (KV-store scatter, split-KV decode, merge).
N > 1  (configs[3], "cfg4"): Llama-3.1-70B shapes KV-head-sharded over N GPUs exactly like SGLang's TP
       (srt/models/llama.py:158-171): per GPU (64/N, 8/N) heads, batch 32 at 32K context, all 80 layers.  One step =
       1 planner launch + 80 x (store, decode, merge, o_proj GEMM fused with its TP all-reduce over NVLink --
       the one collective of the attention block, srt/layers/linear.py:1563-1627).  Total work is fixed as N grows
       ("scaling": "strong").

One JSON line on stdout (rank 0): metric = decode-attention HBM GB/s (algorithmic KV bytes of the step / time,
whole job).  `--impl reference` times the reference's own CPU path (the real TorchNativeAttnBackend imported from
the staged tree oracle/_ref) on the host cores for the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent
if str(REPO) not in sys.path:
    sys.path.insert(0, str(REPO))

import torch  # noqa: E402

METRIC = "decode_attention_hbm_throughput"
UNIT = "GB/s"

# model shapes.  Heads are the FULL model's; a rank gets heads / world.
CFG2 = dict(name="cfg2", model="llama-3.1-8b", layers=32, Hq=32, Hkv=8, D=128, hidden=4096, bs=64, ctx=4096,
            page_size=1, workload="llama-3.1-8b decode attention, bs 64 @ 4096 ctx, page_size 1 (BASELINE cfg2)")
CFG4 = dict(name="cfg4", model="llama-3.1-70b", layers=80, Hq=64, Hkv=8, D=128, hidden=8192, bs=32, ctx=32768,
            page_size=1, workload="llama-3.1-70b decode attention + fused o_proj all-reduce, KV-head-sharded TP, "
                                  "bs 32 @ 32768 ctx, page_size 1 (BASELINE cfg4)")
KV_BUDGET_BYTES = 100e9     # distinct KV layer buffers per GPU (cfg4 at TP2 would need 172 GB for 80)


def decode_bytes(bs, ctx, Hq, Hkv, D, s_kv=2, layers=1):
    """Algorithmic bytes of decode attention: KV rows + page table + q/o."""
    per_layer = bs * ctx * Hkv * 2 * D * s_kv + bs * ctx * 4 + bs * Hq * 2 * D * 2
    return per_layer * layers


def extend_flops(bs, P, E, Hq, D):
    return 4 * Hq * D * bs * (E * P + E * (E + 1) / 2)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "cfg2", "cfg4"],
                    help="auto: cfg2 on one GPU, cfg4 (70B head shards + fused o_proj all-reduce) on N > 1")
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (debugging only)")
    ap.add_argument("--bs", type=int, default=0, help="override the batch size (cfg4 is also reported at bs 8)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the side measurements (cfg5 extend record, cfg3 / cfg4-shard decode, KV store)")
    ap.add_argument("--kernel", default="auto", choices=["auto", "generic"])
    ap.add_argument("--no-graph", action="store_true",
                    help="time eager launches instead of replaying the step as one CUDA graph")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "50",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self, t0: float, t1: float):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        rows = [l.split(", ") for t, l in self.lines if t0 <= t <= t1 + 0.06] or [l.split(", ") for _, l in self.lines]
        sm, reasons, mx, pw = [], set(), None, []
        for r in rows:
            try:
                sm.append(float(r[0])); mx = float(r[1]); pw.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if val.strip().lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "power_w_max": max(pw) if pw else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def pick_workload(args, world):
    cfg = dict(CFG4 if (args.workload == "cfg4" or (args.workload == "auto" and world > 1)) else CFG2)
    if args.layers:
        cfg["layers"] = args.layers
    if args.bs:
        cfg["bs"] = args.bs
        cfg["workload"] = cfg["workload"].replace("bs 32", f"bs {args.bs}").replace("bs 64", f"bs {args.bs}")
    return cfg


def shared_config(cfg, world):
    """The `config` object both arms print (identical for the same N)."""
    return {"workload": cfg["workload"], "model": cfg["model"], "layers": cfg["layers"], "batch": cfg["bs"],
            "context": cfg["ctx"], "heads": [cfg["Hq"], cfg["Hkv"]], "head_dim": cfg["D"],
            "heads_per_gpu": [cfg["Hq"] // world, max(cfg["Hkv"] // world, 1)],
            "parallelism": f"tp{world} (KV-head shards over {world} GPU{'s' if world > 1 else ''})"}


def load_peaks():
    pk = REPO / "MEASURED_PEAKS.json"
    return json.loads(pk.read_text()) if pk.exists() else {}


# ----------------------------------------------------------------------------- CPU side: the reference itself
def run_ref_child(extra, timeout=900):
    """benchmarks/ref_torch_native.py in a child process: the reference's real TorchNativeAttnBackend on the host
    cores (GPU hidden from the child).  Returns its JSON dict, or {"unavailable": why}."""
    env = dict(os.environ, PYTHONDONTWRITEBYTECODE="1", CUDA_VISIBLE_DEVICES="")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_RUN_ID"):
        env.pop(k, None)
    try:
        proc = subprocess.run([sys.executable, str(REPO / "benchmarks" / "ref_torch_native.py"), *map(str, extra)],
                              env=env, capture_output=True, text=True, timeout=timeout)
        return json.loads(proc.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"{type(e).__name__}: {e}"}


def port_decode_seconds(bs, cfg, reps_budget_s):
    """Fallback when oracle/_ref is not staged: the fp32 restatement (oracle/torch_native_ref.decode_ref)."""
    from oracle import torch_native_ref as ref

    g = torch.Generator().manual_seed(0)
    ctx, Hq, Hkv, D = cfg["ctx"], cfg["Hq"], cfg["Hkv"], cfg["D"]
    rows = bs * ctx + 1
    kp = torch.randn(rows, Hkv, D, generator=g).to(torch.bfloat16)
    vp = torch.randn(rows, Hkv, D, generator=g).to(torch.bfloat16)
    rtt = torch.zeros(bs + 1, ctx, dtype=torch.int32)
    rtt[1:] = (torch.randperm(rows - 1, generator=g) + 1).view(bs, ctx).to(torch.int32)
    q = torch.randn(bs, Hq, D, generator=g).to(torch.bfloat16)
    rpi, seq = torch.arange(1, bs + 1), torch.full((bs,), ctx, dtype=torch.int32)
    torch.set_num_threads(os.cpu_count() or 1)
    ref.decode_ref(q[:1], kp, vp, rtt, rpi[:1], seq[:1], D ** -0.5)
    reps, t0 = 0, time.perf_counter()
    while reps < 1 or time.perf_counter() - t0 < reps_budget_s:
        ref.decode_ref(q, kp, vp, rtt, rpi, seq, D ** -0.5)
        reps += 1
    return (time.perf_counter() - t0) / reps, reps


def cpu_decode_sample(cfg, sample_bs, steps, warmup, budget_s):
    """(GB/s, seconds per step, steps run, cores, kind, sample description) of the CPU arm on `sample_bs` requests of
    the workload through ONE layer, full heads (the CPU has no TP)."""
    ref = run_ref_child(["--mode", "decode", "--bs", sample_bs, "--ctx", cfg["ctx"], "--hq", cfg["Hq"],
                         "--hkv", cfg["Hkv"], "--d", cfg["D"], "--steps", steps, "--warmup", warmup,
                         "--budget-s", budget_s])
    nbytes = decode_bytes(sample_bs, cfg["ctx"], cfg["Hq"], cfg["Hkv"], cfg["D"])
    what = (f"{sample_bs} of {cfg['bs']} requests x 1 of {cfg['layers']} layers at ctx {cfg['ctx']}, "
            f"Hq {cfg['Hq']} / Hkv {cfg['Hkv']} / D {cfg['D']}, bf16 pool")
    if "seconds_per_step" in ref:
        dt = ref["seconds_per_step"]
        return (nbytes / dt / 1e9, dt, ref["steps"], ref["cores"], "reference",
                what + f"; {ref['backend']} (bf16 SDPA) imported from oracle/_ref, {ref['steps']} steps")
    dt, reps = port_decode_seconds(sample_bs, cfg, budget_s or 10.0)
    return (nbytes / dt / 1e9, dt, reps, os.cpu_count() or 1, "port",
            what + f"; oracle/torch_native_ref.decode_ref (fp32), {reps} reps; reference tree unavailable: "
            + str(ref.get("unavailable"))[:160])


def run_reference(args):
    """--impl reference: the reference's CPU path on the host cores, same workload / metric / unit.  One step = a
    bounded sample of the workload (stated in `config.sample`), so that K steps finish within minutes."""
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    cfg = pick_workload(args, max(world, args.gpus))
    sample_bs = cfg["bs"] if cfg["name"] == "cfg2" else 2        # cfg4: 2 x 32K x (64, 8) heads = 2.1 GB of KV
    value, dt, steps, cores, kind, sample = cpu_decode_sample(cfg, sample_bs, args.steps, max(args.warmup, 1), 0)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": max(args.warmup, 1), "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True,
        "scaling": "weak" if cfg["name"] == "cfg2" else "strong", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": shared_config(cfg, max(world, args.gpus)), "sample": sample,
        "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------- GPU side
class DecodeStep:
    """Pools, backend, per-layer inputs and (N > 1) the fused o_proj + all-reduce of one decode step."""

    def __init__(self, cfg, device, rank, world, group, seed):
        from sglang_b200.backend import B200AttnBackend
        from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention, make_model_runner
        from sglang_b200.tp import shard_heads

        self.cfg, self.device, self.rank, self.world = cfg, device, rank, world
        shard = shard_heads(cfg["Hq"], cfg["Hkv"], world, rank)
        self.Hq, self.Hkv = shard.q_heads, shard.kv_heads
        bs, ctx, D, L = cfg["bs"], cfg["ctx"], cfg["D"], cfg["layers"]
        Hq, Hkv = self.Hq, self.Hkv
        rows = bs * ctx + cfg["page_size"]
        per_layer = 2 * rows * Hkv * D * 2
        self.n_buf = int(max(1, min(L, KV_BUDGET_BYTES // per_layer)))
        self.runner = make_model_runner(device=device, dtype=torch.bfloat16, pool_dtype=torch.bfloat16, num_heads=Hq,
                                        num_kv_heads=Hkv, head_dim=D, page_size=cfg["page_size"], pool_rows=rows,
                                        max_bs=bs, max_context_len=ctx, layer_num=self.n_buf)
        g = torch.Generator(device=device).manual_seed(seed)
        pool, rtp = self.runner.token_to_kv_pool, self.runner.req_to_token_pool
        pool.k_buffer[0].normal_(generator=g)
        pool.v_buffer[0].normal_(generator=g)
        for l in range(1, self.n_buf):           # distinct buffers (each >> L2); the values need not differ
            pool.k_buffer[l].copy_(pool.v_buffer[l - 1])
            pool.v_buffer[l].copy_(pool.k_buffer[l - 1])
        perm = torch.randperm(rows - 1, generator=g, device=device) + 1
        rtp.req_to_token[1:].copy_(perm[: bs * ctx].view(bs, ctx).to(torch.int32))
        self.backend = B200AttnBackend(self.runner)
        # layer l reads / writes KV buffer l % n_buf: all `layers` launches run, every one streams a buffer far
        # larger than L2 from HBM
        self.layers = [RadixAttention(num_heads=Hq, head_dim=D, scaling=D ** -0.5, num_kv_heads=Hkv,
                                      layer_id=l % self.n_buf) for l in range(L)]
        rpi = torch.arange(1, bs + 1, dtype=torch.int64, device=device)
        seq = torch.full((bs,), ctx, dtype=torch.int32, device=device)
        loc = rtp.req_to_token[1:, ctx - 1].to(torch.int64).contiguous()
        self.fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs, req_pool_indices=rpi, seq_lens=seq,
                               out_cache_loc=loc, seq_lens_sum=bs * ctx)
        self.width = (Hq + 2 * Hkv) * D
        self.qkv = torch.randn(L, bs, self.width, generator=g, device=device, dtype=torch.bfloat16)
        self.comm = self.w = None
        if world > 1:
            from sglang_b200.comm import OProjAllReduce

            # o_proj weight shard of every layer [hidden, Hq_local * D] (RowParallelLinear, linear.py:1563-1627)
            self.w = (torch.randn(L, cfg["hidden"], Hq * D, generator=g, device=device, dtype=torch.float32)
                      * 0.02).to(torch.bfloat16)
            self.comm = OProjAllReduce(group, bs, cfg["hidden"], torch.bfloat16, device)
        self.out_width = cfg["hidden"] if world > 1 else Hq * D

    def split(self, l):
        Hq, Hkv, D = self.Hq, self.Hkv, self.cfg["D"]
        x = self.qkv[l]
        return x[:, : Hq * D], x[:, Hq * D : (Hq + Hkv) * D], x[:, (Hq + Hkv) * D :]

    def layer(self, l, fb=None, oproj=True):
        q, k, v = self.split(l)
        o = self.backend.forward_decode(q, k, v, self.layers[l], fb or self.fb)
        if self.comm is not None and oproj:
            o = self.comm.forward(o, self.w[l])
        return o

    def step(self, events=None, oproj=True):
        be = self.backend
        be.init_forward_metadata(self.fb)
        for l in range(self.cfg["layers"]):
            be.attn_events = events[l] if events is not None else None
            self.layer(l, oproj=oproj)
        be.attn_events = None

    def capture(self, layer_range=None, oproj=True, plan=True, keep_outputs=False):
        """One CUDA graph of the step (decode_cuda_graph_runner.py:946-953 capture, :1168 replay)."""
        from sglang_b200 import ops

        be, dev = self.backend, self.device
        rng = layer_range or range(self.cfg["layers"])
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            graph = torch.cuda.CUDAGraph()
            be.init_forward_metadata_out_graph(self.fb, in_capture=True)
            n0, outs = ops.launch_count, []
            with torch.cuda.graph(graph, stream=side):
                if plan:
                    be.init_forward_metadata_in_graph(self.fb)
                for l in rng:
                    o = self.layer(l, oproj=oproj)
                    if keep_outputs:         # the fused o_proj returns a view of its (reused) symmetric buffer
                        outs.append(o.clone() if self.comm is not None and oproj else o)
                out = torch.stack(outs) if keep_outputs else None        # one D2H copy per segment
            launches = ops.launch_count - n0
        torch.cuda.current_stream().wait_stream(side)
        return graph, launches, out


def time_graph(graph, prep, steps, barrier):
    for _ in range(3):
        prep(); graph.replay()
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        prep(); graph.replay()
    b.record()
    barrier()
    return a.elapsed_time(b) / steps


def measure_extend(device, peaks, page=1):
    """BASELINE configs[4] (cfg5): 128 requests sharing a 16K-token prefix, 1K-token extend each, Llama-8B heads;
    one layer's extend attention, CUDA events, own clocks sample.  The second headline of `metric`.
    page 1 is the reference Triton backend's default pool (K/V rows gathered one by one with cp.async); page 64 is
    the page-granular pool the reference's trtllm_mha / fa4 backends require (K/V pages fetched by TMA boxes)."""
    sys.path.insert(0, str(REPO / "benchmarks"))
    import sweep

    bs, P, E, Hq, Hkv, D = 128, 16384, 1024, 32, 8, 128
    sampler = ClockSampler(device.index or 0)
    sampler.start()
    time.sleep(0.15)
    t0 = time.perf_counter()
    r = sweep.extend_case("cfg5: 128 req x (16K shared prefix + 1K new)", bs, P, E, Hq, Hkv, D, 5, page=page)
    torch.cuda.synchronize()
    clocks = sampler.stop(t0, time.perf_counter())
    tf = r["tflops"]
    sustained, burst = peaks.get("bf16_tflops_sustained"), peaks.get("bf16_tflops")
    peak = float(sustained or 1413.7)
    return {
        "metric": "extend_attention_tflops", "value": tf, "unit": "TFLOP/s", "ms": r["ms"],
        "config": {"workload": f"llama-3.1-8b extend attention, 128 req x (16384 shared prefix + 1024 new), page_size {page} "
                               "(BASELINE cfg5)", "layers": 1, "heads": [Hq, Hkv], "head_dim": D,
                   "kv_fill": "tma_page_boxes" if r.get("page_box_rows") else "cp_async_row_gather"},
        "flops": extend_flops(bs, P, E, Hq, D),
        "roofline": {"bound": "tensor", "kernel": "extend_tc_kernel<bf16, D128>", "achieved": tf, "peak": peak,
                     "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (a 30 ms kernel repeated back to back "
                                    "runs under the power cap)" if sustained else "fallback",
                     "unit": "TFLOP/s", "frac": round(tf / peak, 4),
                     "frac_of_measured_burst": round(tf / float(burst), 4) if burst else None,
                     "frac_of_2250_nominal": round(tf / 2250.0, 4), "kernel_ms": r["ms"], "traffic": None,
                     "timing": "CUDA events around each launch, median of 5 after 2 warm-ups"},
        "clocks": clocks,
    }


def secondary_measurements():
    """Other BASELINE configs, one layer each, CUDA-event timed through the same ops API (benchmarks/sweep.py holds
    the case builders).  Reported beside the headline, not part of `value`."""
    sys.path.insert(0, str(REPO / "benchmarks"))
    import sweep

    bf, e4 = torch.bfloat16, torch.float8_e4m3fn
    out = {}
    try:
        r = sweep.decode_case("cfg3 8B fp8-e4m3 KV, bs 256 @ 8192 ctx, page 1", 256, 8192, 32, 8, 128, e4, 1, 10)
        out["cfg3_fp8_decode"] = {"GBps": r["gbs"], "us": r["us"], "frac_of_measured_hbm": r["frac_measured"],
                                  "frac_of_8TBs": r["frac_8tbs"]}
        for tp, (hq, hkv) in ((2, (32, 4)), (4, (16, 2)), (8, (8, 1))):
            for bs in (32, 8):
                r = sweep.decode_case(f"cfg4 70B TP{tp} shard bs {bs} @ 32768", bs, 32768, hq, hkv, 128, bf, 1, 10)
                out[f"cfg4_tp{tp}_bs{bs}_decode_one_gpu"] = {
                    "GBps": r["gbs"], "us": r["us"], "frac_of_measured_hbm": r["frac_measured"],
                    "frac_of_8TBs": r["frac_8tbs"]}
        r = sweep.store_case("cfg5 KV store 131072 rows bf16", 131072, 8, 128, bf, 10)
        out["cfg5_kv_store"] = {"GBps": r["gbs"], "us": r["us"], "frac_of_measured_hbm": r["frac_measured"]}
        # outside BASELINE's configs, same ops API: speculative verify (tree mask over 8 draft tokens) with the
        # backend's split-KV choice, and the head dims / pool dtypes that left the CUDA-core kernels this round
        for bs, ctx in ((1, 32768), (8, 16384), (64, 4096)):
            r0 = sweep.extend_case("verify unsplit", bs, ctx, 8, 32, 8, 128, 5, shared=False, tree_mask=True,
                                   one_tile=True, splits=1)
            r1 = sweep.extend_case("verify split", bs, ctx, 8, 32, 8, 128, 5, shared=False, tree_mask=True,
                                   one_tile=True, splits=0)
            out[f"verify_bs{bs}_ctx{ctx}"] = {"ms": r1["ms"], "ms_unsplit": r0["ms"], "splits": r1["splits"]}
        r = sweep.decode_case("gpt-oss heads 64/8 D 64, fp8-e4m3 KV, bs 64 @ 8192", 64, 8192, 64, 8, 64, e4, 1, 10)
        out["d64_fp8_decode"] = {"GBps": r["gbs"], "us": r["us"], "frac_of_8TBs": r["frac_8tbs"]}
        r = sweep.decode_case("phi-3-mini heads 32/32 D 96, bs 64 @ 8192", 64, 8192, 32, 32, 96, bf, 1, 10)
        out["d96_decode"] = {"GBps": r["gbs"], "us": r["us"], "frac_of_8TBs": r["frac_8tbs"]}
    except Exception as e:  # side measurements must never take the headline line down
        out["error"] = f"{type(e).__name__}: {e}"
    return out


def run_ours(args):
    rank, world, local = dist_env()
    import torch.distributed as dist

    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
        group = dist.group.WORLD
    from sglang_b200 import ops

    cfg = pick_workload(args, world)
    layers, bs = cfg["layers"], cfg["bs"]
    st = DecodeStep(cfg, device, rank, world, group, seed=rank)
    be = st.backend
    if args.kernel == "generic":
        be.decode_kernel = ops.KERNEL_GENERIC
    Hq, Hkv, D = st.Hq, st.Hkv, cfg["D"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        st.step()
    barrier()

    graph, per_graph, ms_no_oproj = None, 0, None
    prep = lambda: be.init_forward_metadata_out_graph(st.fb)   # noqa: E731  (the runner's replay prep)
    if not args.no_graph:
        be.init_cuda_graph_state(bs, bs)
        st.step()
        graph, per_graph, _ = st.capture()
        if st.comm is not None:                    # the same step without the fused o_proj: its share of the step
            g2, _, _ = st.capture(oproj=False)
            ms_no_oproj = time_graph(g2, prep, max(args.steps // 4, 5), barrier)
            del g2
        for _ in range(3):
            prep(); graph.replay()

    # live per-kernel timing: one event pair per (step, layer) around the decode attention(+merge) launch
    ev_steps = args.steps if graph is None else min(args.steps, max(2, 640 // layers))
    pairs = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(layers)]
             for _ in range(ev_steps)]
    for s_ in pairs:
        for a, b in s_:
            a.record(); b.record()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.15)
    barrier()
    n0 = ops.launch_count
    ev0, ev1, ev2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    t0 = time.perf_counter()
    ev0.record()
    if graph is not None:
        for _ in range(args.steps):
            prep(); graph.replay()
    ev1.record()
    # eager launches of the same step with CUDA events around every attention(+merge) launch: the whole timed region
    # with --no-graph, else a shorter instrumented pass (events cannot bracket kernels inside a replay)
    for s in range(ev_steps):
        st.step(pairs[s])
    ev2.record()
    barrier()
    t1 = time.perf_counter()
    launches = ops.launch_count - n0 + (per_graph * args.steps if graph is not None else 0)
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    eager_ms = ev1.elapsed_time(ev2) / ev_steps
    ms = ev0.elapsed_time(ev1) / args.steps if graph is not None else eager_ms
    kern_ms = sorted(a.elapsed_time(b) for s_ in pairs for a, b in s_)
    kern_avg = sum(kern_ms) / len(kern_ms)

    # ---- e2e: the same step through the backend API with HOST buffers (pinned), copies inside the timed region.
    # Copies run on their own streams so that the PCIe transfers of the next layers' inputs and the previous layers'
    # outputs overlap the kernels; every byte still crosses inside the timed region, every step, and the step ends
    # when the last output byte is on the host.  With graphs the step is cut into segments, one CUDA graph each.
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode

    fb = st.fb
    h_qkv = torch.randn(layers, bs, st.width).to(torch.bfloat16).pin_memory()
    h_meta = torch.stack([fb.req_pool_indices.cpu(), fb.seq_lens.cpu().long(), fb.out_cache_loc.cpu()]).pin_memory()
    h_out = torch.empty(layers, bs, st.out_width, dtype=torch.bfloat16).pin_memory()
    d_meta = torch.empty_like(h_meta, device=device)
    s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
    seg = 1
    if graph is not None:
        seg = next(c for c in (8, 4, 2, 1) if layers % c == 0) if layers > 32 else (4 if layers % 4 == 0 else 1)
    n_seg = layers // seg
    in_ev = [torch.cuda.Event() for _ in range(n_seg)]
    done_ev = [torch.cuda.Event() for _ in range(n_seg)]
    seg_graphs, seg_outs = [], []
    if graph is not None:
        for i in range(n_seg):
            g_i, _, out_i = st.capture(range(i * seg, (i + 1) * seg), plan=(i == 0), keep_outputs=True)
            seg_graphs.append(g_i)
            seg_outs.append(out_i)

    def e2e_step():
        cur = torch.cuda.current_stream()
        s_in.wait_stream(cur)                      # previous step's readers of qkv are done (the step ends in a sync)
        with torch.cuda.stream(s_in):
            for i in range(n_seg):
                st.qkv[i * seg:(i + 1) * seg].copy_(h_qkv[i * seg:(i + 1) * seg], non_blocking=True)
                in_ev[i].record(s_in)
        if graph is not None:
            # the runner's static metadata buffers are refreshed from the host, then the replay prep
            fb.req_pool_indices.copy_(h_meta[0], non_blocking=True)
            fb.seq_lens.copy_(h_meta[1], non_blocking=True)
            fb.out_cache_loc.copy_(h_meta[2], non_blocking=True)
            be.init_forward_metadata_out_graph(fb)
            fb2 = fb
        else:
            d_meta.copy_(h_meta, non_blocking=True)
            fb2 = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs, req_pool_indices=d_meta[0],
                               seq_lens=d_meta[1], out_cache_loc=d_meta[2], seq_lens_sum=bs * cfg["ctx"])
            be.init_forward_metadata(fb2)
        for i in range(n_seg):
            cur.wait_event(in_ev[i])
            if graph is not None:
                seg_graphs[i].replay()
                o = seg_outs[i]
            else:
                o = st.layer(i, fb2).unsqueeze(0)
                o.record_stream(s_out)
            done_ev[i].record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(done_ev[i])
                h_out[i * seg:(i + 1) * seg].copy_(o.view(seg, bs, -1), non_blocking=True)
        s_out.synchronize()                        # the caller reads the result
        cur.synchronize()

    for _ in range(3):
        e2e_step()
    barrier()
    e2e_steps = max(args.steps // 2, 5)
    te0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - te0) / e2e_steps * 1e3
    # the bytes that reached the host are the outputs of the inputs that were copied in: recompute two layers on the
    # device -- for N > 1 through cuBLAS + an NCCL all-reduce, so the fused o_proj kernel's exchange is checked too
    check = {}
    for l in (0, layers - 1):
        q, k, v = st.split(l)
        want = be.forward_decode(q, k, v, st.layers[l], fb).float()
        if st.comm is not None:
            want = (want.to(torch.bfloat16) @ st.w[l].t()).float()
            dist.all_reduce(want)
        got = h_out[l].float().to(device).view_as(want)
        rel = float((got - want).pow(2).mean().sqrt() / want.pow(2).mean().sqrt().clamp_min(1e-20))
        err = float((got - want).abs().max())
        check[f"layer{l}"] = {"max_abs": round(err, 5), "rel_rms": round(rel, 5)}
        if not (rel <= 2e-2 and (st.comm is not None or err <= 2e-2)):
            raise RuntimeError(f"e2e output of layer {l} differs from the device-side result: {check}")
    h2d = h_qkv.numel() * 2 + h_meta.numel() * 8
    d2h = h_out.numel() * 2

    t = torch.tensor([ms, e2e_ms, kern_avg, eager_ms, ms_no_oproj or 0.0], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms, kern_avg, eager_ms, ms_no_oproj = (float(x) for x in t.cpu())
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    per_gpu_bytes = decode_bytes(bs, cfg["ctx"], Hq, Hkv, D, layers=layers)
    step_bytes = per_gpu_bytes * world
    value = step_bytes / (ms * 1e-3) / 1e9
    e2e_value = step_bytes / (e2e_ms * 1e-3) / 1e9
    peaks = load_peaks()
    peak, peak_src = (float(peaks["hbm_gbs"]), "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    achieved = decode_bytes(bs, cfg["ctx"], Hq, Hkv, D) / (kern_avg * 1e-3) / 1e9
    traffic = None
    tf = REPO / "profiles" / "decode_traffic.json"
    if tf.exists() and cfg["name"] == "cfg2":
        traffic = json.loads(tf.read_text()).get("dram_bytes_per_launch")
    strong = cfg["name"] == "cfg4"
    kernel_name = ("decode_hg_kernel<128,bf16>" if Hkv % 4 == 0 else "decode_mma_kernel<128,bf16>") + \
        " + decode_merge_fast_kernel"
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(ms, 4), "higher_is_better": True,
        "scaling": "strong" if strong else "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": shared_config(cfg, world),
        "run": {"collective": ("o_proj GEMM fused with its all-reduce (one kernel per layer, NVLink/NVSwitch peer "
                               f"memory, multicast={'on' if st.comm.multicast else 'off'})") if st.comm else "none (TP1)",
                "kv_layer_buffers": st.n_buf if st.n_buf == layers else
                f"{st.n_buf} distinct buffers cycled over {layers} layer launches ({layers} x "
                f"{per_gpu_bytes / layers / 1e9:.2f} GB exceeds HBM at this TP; every buffer is >> L2)",
                "l2": f"inputs larger than L2 ({per_gpu_bytes / layers / 1e9:.2f} GB of KV per layer launch vs "
                      "126 MB L2); no flush",
                "decode_kernel": args.kernel,
                "launch": ("one CUDA graph of the whole step, replayed after init_forward_metadata_out_graph "
                           "(how SGLang's decode runner drives a backend)" if graph is not None else "eager launches")},
        "ms_per_step_eager": round(eager_ms, 4),
        "tokens_per_s": round(bs / (ms * 1e-3), 1) if strong else round(bs * world / (ms * 1e-3), 1),
        "per_gpu_GBps": round(value / world, 1),
        "pct_of_8TBs": round(100 * value / world / 8000.0, 2),
        "roofline": {"bound": "hbm", "kernel": kernel_name, "achieved": round(achieved, 1), "peak": peak,
                     "peak_source": peak_src, "unit": "GB/s", "frac": round(achieved / peak, 4),
                     "frac_of_8TBs_nominal": round(achieved / 8000.0, 4), "kernel_us_avg": round(kern_avg * 1e3, 2),
                     "kernel_us_median": round(kern_ms[len(kern_ms) // 2] * 1e3, 2), "traffic": traffic,
                     "timing": f"CUDA events around each attention+merge launch pair, {ev_steps} eager steps x {layers} "
                               "layers inside the timed run",
                     "peak_note": "peak is the measured COPY bandwidth (read + write); this kernel only reads, and a "
                                  "read-only stream of >= 512-byte rows reaches 6.7-7.3 TB/s on this part "
                                  "(profiles/r1_gather_rate.txt), so frac can pass 1.0 on the 1 KiB-row shards"},
        "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "ms_per_step": round(e2e_ms, 4),
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "output_check": check},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    if st.comm is not None and ms_no_oproj:
        line["oproj_allreduce"] = {
            "ms_per_step_without": round(ms_no_oproj, 4), "share_of_step": round(1.0 - ms_no_oproj / ms, 4),
            "us_per_layer": round((ms - ms_no_oproj) / layers * 1e3, 2),
            "shape": f"x [{bs}, {Hq * D}] @ w [{cfg['hidden']}, {Hq * D}]^T, summed over {world} ranks"}
    if world == 1 and not args.no_secondary:
        del st.qkv
        try:
            line["extend"] = measure_extend(device, peaks)
            line["extend_page64"] = measure_extend(device, peaks, page=64)
        except Exception as e:  # noqa: BLE001
            line.setdefault("extend", {"error": f"{type(e).__name__}: {e}"})
            line.setdefault("extend_page64", {"error": f"{type(e).__name__}: {e}"})
        line["secondary"] = secondary_measurements()
    if not args.no_cpu_baseline and world == 1:
        sample_bs = 8 if cfg["name"] == "cfg2" else 1
        gbs, dt, reps, cores, kind, sample = cpu_decode_sample(cfg, sample_bs, 1000, 1, 12.0)
        line["cpu_baseline"] = {"value": round(gbs, 3), "unit": UNIT, "cores": cores, "kind": kind,
                                "sample": sample + f", {dt:.3f} s each"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def _reserve_stdout_for_the_json_line():
    """Libraries write to fd 1 behind Python's back (NCCL prints "NCCL version ..." there under torchrun): point fd 1 at
    stderr and keep the real stdout for this script's own print() -- the one JSON line."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)


def main():
    args = parse_args()
    _reserve_stdout_for_the_json_line()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
