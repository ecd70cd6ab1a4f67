#!/usr/bin/env python
"""bench.py -- RadixAttention hot-path benchmark on B200 (contract: see DESIGN.md section "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload at N=1 (BASELINE.json configs[1]): Llama-3.1-8B shapes, bf16, TP1, paged decode,
batch 64 at 4K context, page_size 1, random slot permutation.  One step = one decode step of the
attention path of all 32 layers: 1 planner launch + 32 x (KV-store scatter, split-KV decode, merge).
N>1: every rank runs the same per-rank shape on its own KV-head shard (weak scaling, no collective
inside attention).

One JSON line on stdout (rank 0): metric = decode-attention HBM GB/s (algorithmic bytes / time).
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

# Llama-3.1-8B attention shapes
CFG2 = dict(layers=32, Hq=32, Hkv=8, D=128, bs=64, ctx=4096, page_size=1)


def decode_bytes(bs, ctx, Hq, Hkv, D, s_kv=2, layers=1):
    """Algorithmic bytes of decode attention: KV rows + page table + q/o."""
    per_layer = bs * ctx * Hkv * 2 * D * s_kv + bs * ctx * 4 + bs * Hq * 2 * D * 2
    return per_layer * layers


def store_bytes(T, Hkv, D, s_kv=2, layers=1):
    return (T * Hkv * 2 * D * (2 + s_kv) + T * 8) * layers


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--layers", type=int, default=CFG2["layers"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the one-layer cfg3 (FP8 decode) / cfg5 (prefix-reuse extend) / KV-store side measurements")
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
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ----------------------------------------------------------------------------- CPU side (oracle port)
def cpu_problem(bs, ctx, Hq, Hkv, D, seed=0):
    g = torch.Generator().manual_seed(seed)
    rows = bs * ctx + 1
    kp = torch.randn(rows, Hkv, D, generator=g).to(torch.bfloat16)
    vp = torch.randn(rows, Hkv, D, generator=g).to(torch.bfloat16)
    rtt = torch.zeros(bs + 1, ctx, dtype=torch.int32)
    rtt[1:] = (torch.randperm(rows - 1, generator=g) + 1).view(bs, ctx).to(torch.int32)
    q = torch.randn(bs, Hq, D, generator=g).to(torch.bfloat16)
    return q, kp, vp, rtt, torch.arange(1, bs + 1), torch.full((bs,), ctx, dtype=torch.int32)


def cpu_decode_throughput(sample_bs, cfg, budget_s, min_reps=1, max_reps=1000):
    """Times the oracle restatement of the reference's torch_native decode (its CPU path) on all
    host cores, on `sample_bs` requests x 1 layer of the bench workload.  Returns (GB/s, seconds/rep, reps)."""
    from oracle import torch_native_ref as ref

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    q, kp, vp, rtt, rpi, seq = cpu_problem(sample_bs, cfg["ctx"], cfg["Hq"], cfg["Hkv"], cfg["D"])
    scale = cfg["D"] ** -0.5
    ref.decode_ref(q[:1], kp, vp, rtt, rpi[:1], seq[:1], scale)  # warm-up
    reps, t0 = 0, time.perf_counter()
    while reps < min_reps or (time.perf_counter() - t0 < budget_s and reps < max_reps):
        ref.decode_ref(q, kp, vp, rtt, rpi, seq, scale)
        reps += 1
    dt = (time.perf_counter() - t0) / reps
    gbs = decode_bytes(sample_bs, cfg["ctx"], cfg["Hq"], cfg["Hkv"], cfg["D"]) / dt / 1e9
    return gbs, dt, reps, cores


def run_reference(args):
    """--impl reference: the reference's CPU path (torch_native semantics; oracle port because the
    reference tree does not exist on the GPU box) on the host cores, same config/metric/unit.
    One step = the full batch (64 requests x 4K context) through ONE layer: a bounded sample of the
    32-layer step, so that K steps finish within minutes."""
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    from oracle import torch_native_ref as ref

    cfg = dict(CFG2)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    q, kp, vp, rtt, rpi, seq = cpu_problem(cfg["bs"], cfg["ctx"], cfg["Hq"], cfg["Hkv"], cfg["D"])
    scale = cfg["D"] ** -0.5
    for _ in range(max(args.warmup, 1)):
        ref.decode_ref(q[:8], kp, vp, rtt, rpi[:8], seq[:8], scale)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ref.decode_ref(q, kp, vp, rtt, rpi, seq, scale)
    dt = (time.perf_counter() - t0) / args.steps
    value = decode_bytes(cfg["bs"], cfg["ctx"], cfg["Hq"], cfg["Hkv"], cfg["D"]) / dt / 1e9
    sample = "1 of 32 layers per step: bs 64 x ctx 4096, Hq 32 / Hkv 8 / D 128, bf16 pool, fp32 math"
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "llama-3.1-8b decode attention, bs 64 @ 4096 ctx, page_size 1 (BASELINE cfg2)",
                   "sample": sample},
        "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------- GPU side
def build_workload(cfg, layers, device, seed):
    """Mirror pools + backend + per-layer inputs for the decode step (everything resident in HBM)."""
    sys.path.insert(0, str(REPO / "tests"))
    from helpers import make_runner

    from sglang_b200.backend import B200AttnBackend
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention

    bs, ctx, Hq, Hkv, D = cfg["bs"], cfg["ctx"], cfg["Hq"], cfg["Hkv"], cfg["D"]
    rows = bs * ctx + cfg["page_size"]
    runner = make_runner(device=device, dtype=torch.bfloat16, pool_dtype=torch.bfloat16, Hq=Hq, Hkv=Hkv, D=D,
                         page_size=cfg["page_size"], pool_rows=rows, max_bs=bs, max_context_len=ctx,
                         layer_num=layers)
    g = torch.Generator(device=device).manual_seed(seed)
    pool, rtp = runner.token_to_kv_pool, runner.req_to_token_pool
    for l in range(layers):
        pool.k_buffer[l].normal_(generator=g)
        pool.v_buffer[l].normal_(generator=g)
    perm = torch.randperm(rows - 1, generator=g, device=device) + 1
    rtp.req_to_token[1:].copy_(perm[: bs * ctx].view(bs, ctx).to(torch.int32))
    backend = B200AttnBackend(runner)
    layer_objs = [RadixAttention(num_heads=Hq, head_dim=D, scaling=D**-0.5, num_kv_heads=Hkv, layer_id=l)
                  for l in range(layers)]
    rpi = torch.arange(1, bs + 1, dtype=torch.int64, device=device)
    seq = torch.full((bs,), ctx, dtype=torch.int32, device=device)
    loc = rtp.req_to_token[1:, ctx - 1].to(torch.int64).contiguous()
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs, req_pool_indices=rpi, seq_lens=seq,
                      out_cache_loc=loc, seq_lens_sum=bs * ctx)
    width = (Hq + 2 * Hkv) * D
    qkv = torch.randn(layers, bs, width, generator=g, device=device, dtype=torch.bfloat16)
    return runner, backend, layer_objs, fb, qkv


def split_qkv(qkv_l, cfg):
    Hq, Hkv, D = cfg["Hq"], cfg["Hkv"], cfg["D"]
    return qkv_l[:, : Hq * D], qkv_l[:, Hq * D : (Hq + Hkv) * D], qkv_l[:, (Hq + Hkv) * D :]


def secondary_measurements(*release):
    """The other BASELINE.json configs, one layer each, CUDA-event timed through the same ops API
    (benchmarks/sweep.py holds the case builders).  Reported beside the headline, not part of `value`."""
    sys.path.insert(0, str(REPO / "benchmarks"))
    import sweep

    bf, e4 = torch.bfloat16, torch.float8_e4m3fn
    out = {}
    try:
        r = sweep.decode_case("cfg3 8B fp8-e4m3 KV, bs 256 @ 8192 ctx, page 1", 256, 8192, 32, 8, 128, e4, 1, 10)
        out["cfg3_fp8_decode"] = {"GBps": r["gbs"], "us": r["us"], "frac_of_measured_hbm": r["frac_measured"],
                                  "frac_of_8TBs": r["frac_8tbs"]}
        r = sweep.decode_case("cfg4 70B TP8 shard (Hq 8 / Hkv 1), bs 32 @ 32768 ctx", 32, 32768, 8, 1, 128, bf, 1, 10)
        out["cfg4_tp8_decode"] = {"GBps": r["gbs"], "us": r["us"], "frac_of_measured_hbm": r["frac_measured"],
                                  "frac_of_8TBs": r["frac_8tbs"]}
        r = sweep.extend_case("cfg5: 128 req x (16K shared prefix + 1K new)", 128, 16384, 1024, 32, 8, 128, 3)
        out["cfg5_extend"] = {"TFLOPs": r["tflops"], "ms": r["ms"], "frac_of_measured_bf16": r["frac_measured"],
                              "frac_of_2250_nominal": r["frac_nominal"]}
        r = sweep.store_case("cfg5 KV store 131072 rows bf16", 131072, 8, 128, bf, 10)
        out["cfg5_kv_store"] = {"GBps": r["gbs"], "us": r["us"], "frac_of_measured_hbm": r["frac_measured"]}
    except Exception as e:  # side measurements must never take the headline line down
        out["error"] = f"{type(e).__name__}: {e}"
    return out


def run_ours(args):
    rank, world, local = dist_env()
    import torch.distributed as dist

    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    from sglang_b200 import ops

    cfg, layers = dict(CFG2), args.layers
    runner, backend, layer_objs, fb, qkv = build_workload(cfg, layers, device, seed=rank)
    if args.kernel == "generic":
        backend.decode_kernel = ops.KERNEL_GENERIC
    bs, Hq, Hkv, D = cfg["bs"], cfg["Hq"], cfg["Hkv"], cfg["D"]

    def step(events=None):
        backend.init_forward_metadata(fb)
        for l in range(layers):
            q, k, v = split_qkv(qkv[l], cfg)
            backend.attn_events = events[l] if events is not None else None
            backend.forward_decode(q, k, v, layer_objs[l], fb)
        backend.attn_events = None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()

    # ---- the step as the serving engine runs it: one CUDA graph of the whole 32-layer decode step, replayed after
    # the backend's replay prep (decode_cuda_graph_runner.py:946-953 capture, :1168 replay)
    graph, per_graph = None, 0
    if not args.no_graph:
        backend.init_cuda_graph_state(bs, bs)
        side = torch.cuda.Stream(device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            step()                                   # warm-up on the capture stream with the graph-state workspaces
            graph = torch.cuda.CUDAGraph()
            backend.init_forward_metadata_out_graph(fb, in_capture=True)
            n_cap = ops.launch_count
            graph_outs = []
            with torch.cuda.graph(graph, stream=side):
                backend.init_forward_metadata_in_graph(fb)
                for l in range(layers):
                    q, k, v = split_qkv(qkv[l], cfg)
                    graph_outs.append(backend.forward_decode(q, k, v, layer_objs[l], fb))
            per_graph = ops.launch_count - n_cap
        torch.cuda.current_stream().wait_stream(side)

        def graph_step():
            backend.init_forward_metadata_out_graph(fb)
            graph.replay()

        for _ in range(3):
            graph_step()

    # live per-kernel timing: one event pair per (step, layer) around the decode attention kernel
    ev_steps = args.steps if graph is None else min(args.steps, 20)
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
        for s in range(args.steps):
            graph_step()
    ev1.record()
    # eager launches of the same step with CUDA events around every attention(+merge) launch: the whole timed
    # region when --no-graph, else a shorter instrumented pass (events cannot bracket kernels inside a replay)
    for s in range(ev_steps):
        step(pairs[s])
    ev2.record()
    barrier()
    t1 = time.perf_counter()
    launches = ops.launch_count - n0 + (per_graph * args.steps if graph is not None else 0)
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    eager_ms = ev1.elapsed_time(ev2) / ev_steps
    ms = ev0.elapsed_time(ev1) / args.steps if graph is not None else eager_ms
    kern_ms = sorted(a.elapsed_time(b) for s_ in pairs for a, b in s_)
    kern_avg = sum(kern_ms) / len(kern_ms)

    # ---- e2e: same step through the backend API with HOST buffers (pinned), copies inside the timed region
    width = (Hq + 2 * Hkv) * D
    h_qkv = torch.randn(layers, bs, width).to(torch.bfloat16).pin_memory()
    h_meta = torch.stack([fb.req_pool_indices.cpu(), fb.seq_lens.cpu().long(), fb.out_cache_loc.cpu()]).pin_memory()
    h_out = torch.empty(layers, bs, Hq * D, dtype=torch.bfloat16).pin_memory()
    d_meta = torch.empty_like(h_meta, device=device)
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode

    # Copies run on their own streams so that the PCIe transfers of the next layers' inputs and the previous layers'
    # outputs overlap the kernels; every byte still crosses inside the timed region, every step, and the step ends
    # when the last output byte is on the host.  With graphs the step is cut into segments of `seg` layers, one
    # CUDA graph each (driving 96 launches + 64 copies per step from Python is CPU-bound otherwise).
    s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
    seg = 4 if (graph is not None and layers % 4 == 0) else 1
    n_seg = layers // seg
    in_ev = [torch.cuda.Event() for _ in range(n_seg)]
    done_ev = [torch.cuda.Event() for _ in range(n_seg)]
    seg_graphs, seg_outs = [], []
    if graph is not None:
        side = torch.cuda.Stream(device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            backend.init_forward_metadata_out_graph(fb, in_capture=True)
            for i in range(n_seg):
                g_i, outs = torch.cuda.CUDAGraph(), []
                with torch.cuda.graph(g_i, stream=side):
                    if i == 0:
                        backend.init_forward_metadata_in_graph(fb)
                    for l in range(i * seg, (i + 1) * seg):
                        q, k, v = split_qkv(qkv[l], cfg)
                        outs.append(backend.forward_decode(q, k, v, layer_objs[l], fb))
                    out_i = torch.stack(outs)            # one D2H copy per segment
                seg_graphs.append(g_i)
                seg_outs.append(out_i)
        torch.cuda.current_stream().wait_stream(side)

    def e2e_step():
        cur = torch.cuda.current_stream()
        s_in.wait_stream(cur)                      # previous step's readers of qkv are done (the step ends in a sync)
        with torch.cuda.stream(s_in):
            for i in range(n_seg):
                qkv[i * seg:(i + 1) * seg].copy_(h_qkv[i * seg:(i + 1) * seg], non_blocking=True)
                in_ev[i].record(s_in)
        if graph is not None:
            # the runner's static metadata buffers are refreshed from the host, then the replay prep
            fb.req_pool_indices.copy_(h_meta[0], non_blocking=True)
            fb.seq_lens.copy_(h_meta[1], non_blocking=True)
            fb.out_cache_loc.copy_(h_meta[2], non_blocking=True)
            backend.init_forward_metadata_out_graph(fb)
            fb2 = fb
        else:
            d_meta.copy_(h_meta, non_blocking=True)
            fb2 = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs, req_pool_indices=d_meta[0],
                               seq_lens=d_meta[1], out_cache_loc=d_meta[2], seq_lens_sum=bs * cfg["ctx"])
            backend.init_forward_metadata(fb2)
        for i in range(n_seg):
            cur.wait_event(in_ev[i])
            if graph is not None:
                seg_graphs[i].replay()
                o = seg_outs[i]
            else:
                q, k, v = split_qkv(qkv[i], cfg)
                o = backend.forward_decode(q, k, v, layer_objs[i], fb2).unsqueeze(0)
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
    # the bytes that reached the host are the attention outputs of the inputs that were copied in
    for l in (0, layers - 1):
        q, k, v = split_qkv(qkv[l], cfg)
        want = backend.forward_decode(q, k, v, layer_objs[l], fb).float().cpu()
        err = float((h_out[l].float() - want.view_as(h_out[l])).abs().max())
        if not err <= 2e-2:
            raise RuntimeError(f"e2e output of layer {l} differs from the device-side result by {err}")
    h2d = h_qkv.numel() * 2 + h_meta.numel() * 8
    d2h = h_out.numel() * 2

    t = torch.tensor([ms, e2e_ms, kern_avg, eager_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms, kern_avg, eager_ms = (float(x) for x in t.cpu())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    step_bytes = decode_bytes(bs, cfg["ctx"], Hq, Hkv, D, layers=layers) * world
    value = step_bytes / (ms * 1e-3) / 1e9
    e2e_value = step_bytes / (e2e_ms * 1e-3) / 1e9
    peaks = {}
    pk = REPO / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    peak, peak_src = (float(peaks["hbm_gbs"]), "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    achieved = decode_bytes(bs, cfg["ctx"], Hq, Hkv, D) / (kern_avg * 1e-3) / 1e9
    traffic = None
    tf = REPO / "profiles" / "decode_traffic.json"
    if tf.exists():
        traffic = json.loads(tf.read_text()).get("dram_bytes_per_launch")
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "llama-3.1-8b decode attention, bs 64 @ 4096 ctx, page_size 1 (BASELINE cfg2)",
                   "layers": layers, "heads_per_gpu": [Hq, Hkv], "head_dim": D, "parallelism": f"tp{world}-head-shards",
                   "l2": "inputs larger than L2 (1.07 GB of KV per layer vs 126 MB L2); no flush",
                   "decode_kernel": args.kernel,
                   "launch": ("one CUDA graph of the whole step, replayed after init_forward_metadata_out_graph "
                              "(how SGLang's decode runner drives a backend)" if graph is not None else "eager launches")},
        "ms_per_step_eager": round(eager_ms, 4),
        "tokens_per_s": round(bs * world / (ms * 1e-3), 1),
        "pct_of_8TBs": round(100 * value / world / 8000.0, 2),
        "roofline": {"bound": "hbm", "kernel": "decode_hg_kernel<128,bf16> + decode_merge_fast_kernel", "achieved": round(achieved, 1), "peak": peak,
                     "peak_source": peak_src, "unit": "GB/s", "frac": round(achieved / peak, 4),
                     "frac_of_8TBs_nominal": round(achieved / 8000.0, 4), "kernel_us_avg": round(kern_avg * 1e3, 2),
                     "kernel_us_median": round(kern_ms[len(kern_ms) // 2] * 1e3, 2), "traffic": traffic,
                     "timing": f"CUDA events around each attention+merge launch pair, {ev_steps} eager steps x {layers} "
                               "layers inside the timed run"},
        "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "ms_per_step": round(e2e_ms, 4),
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    if world == 1 and not args.no_secondary:
        line["secondary"] = secondary_measurements(runner, backend, qkv)
    if not args.no_cpu_baseline:
        gbs, dt, reps, cores = cpu_decode_throughput(8, cfg, budget_s=12.0)
        line["cpu_baseline"] = {"value": round(gbs, 3), "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"8 of 64 requests x 1 of {layers} layers at ctx 4096, {reps} reps, "
                                          f"{dt:.3f} s each (oracle/torch_native_ref.decode_ref, fp32)"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def _reserve_stdout_for_the_json_line():
    """Libraries write to fd 1 behind Python's back (NCCL prints "NCCL version ..." there under torchrun): point fd 1 at
    stderr and keep the real stdout for this script's own print() — the one JSON line."""
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
