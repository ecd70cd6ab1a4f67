#!/usr/bin/env python
"""Kernel-level sweep over BASELINE.json's configs (one layer each), CUDA-event timed.

    python benchmarks/sweep.py [--only decode|extend|store] [--iters 30] [--json out.json]

Prints one line per case: kernel time, achieved GB/s (decode/store) or TFLOP/s (extend) and the
fraction of the measured / nominal B200 peaks.  Not the contract benchmark (that is bench.py);
this is the table DESIGN.md and profiles/ quote for cfg3-cfg5.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
from sglang_b200 import ops  # noqa: E402

PEAKS = json.loads((REPO / "MEASURED_PEAKS.json").read_text()) if (REPO / "MEASURED_PEAKS.json").exists() else {}
HBM = float(PEAKS.get("hbm_gbs", 6650.0))
TF = float(PEAKS.get("bf16_tflops", 1590.0))


def time_cuda(fn, iters, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2] * 1e-3


def time_cuda_graph(fn, reps=20, iters=10):
    """GPU-side time of one call: `reps` calls captured in a CUDA graph and replayed (what a serving
    engine does for decode) -- takes the Python/ctypes launch cost (~25 us per call) out of the number."""
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
    return time_cuda(g.replay, iters, warmup=2) / reps


def decode_case(name, bs, L, Hq, Hkv, D, pool_dtype, page, iters, ragged=False, kernel=0):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    lens = torch.full((bs,), L, dtype=torch.int32, device=dev)
    if ragged:
        lens = torch.randint(L // 2, L + 1, (bs,), generator=g, device=dev, dtype=torch.int32)
    n_pages = bs * ((L + page - 1) // page)
    rows = (n_pages + 1) * page
    es = 1 if pool_dtype == torch.float8_e4m3fn else 2
    kp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    vp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    if es == 1:
        kp, vp = kp.to(pool_dtype).view(torch.uint8), vp.to(pool_dtype).view(torch.uint8)
    perm = torch.randperm(n_pages, generator=g, device=dev) + 1
    ppr = (L + page - 1) // page
    rtt = torch.zeros(bs + 1, ppr * page, dtype=torch.int32, device=dev)
    rtt[1:] = (perm.view(bs, ppr, 1) * page + torch.arange(page, device=dev).view(1, 1, page)).view(bs, -1).int()
    q = torch.randn(bs, Hq, D, generator=g, device=dev, dtype=torch.bfloat16)
    o = torch.empty_like(q)
    rpi = torch.arange(1, bs + 1, device=dev)
    sm, _, _ = ops.device_info()
    ws = ops.DecodeWorkspace.create(bs, Hq, D, dev, num_ctas=sm * ops.decode_ctas_per_sm(D, pool_dtype))
    ops.decode_plan(ws, lens, Hkv, items_per_chunk=ops.decode_items_per_chunk(Hq, Hkv, D, D, torch.bfloat16, pool_dtype))

    def fn():
        ops.decode_attention(q, o, kp, vp, rtt, rpi, lens, ws, D**-0.5, kv_dtype=pool_dtype, kernel=kernel)

    t_eager = time_cuda(fn, iters)
    t, timing = t_eager, "eager"
    # The two launches (attention + merge) cost ~25 us of Python/ctypes time each, which a sub-millisecond kernel
    # pair does not hide (cfg4 TP8: 98.6 us eager vs 88.4 us replayed).  Decode runs under CUDA graphs in the
    # serving engine, so the replay time is the one reported; the eager time stays beside it.
    if t_eager < 500e-6:
        t, timing = time_cuda_graph(fn), "cuda-graph replay"
    tot = int(lens.sum())
    nbytes = tot * Hkv * 2 * D * es + tot * 4 + bs * Hq * 2 * D * 2
    gbs = nbytes / t / 1e9
    return dict(case=name, kind="decode", us=round(t * 1e6, 1), gbs=round(gbs, 1), frac_measured=round(gbs / HBM, 3),
                frac_8tbs=round(gbs / 8000, 3), chunk=int(ws.plan[0]), chunks=int(ws.plan[1]), timing=timing,
                us_eager=round(t_eager * 1e6, 1))


def extend_case(name, bs, P, E, Hq, Hkv, D, iters, shared=True, kernel=0, pool_dtype=torch.bfloat16, tree_mask=False,
                page=1, one_tile=False, splits=1):
    """`splits`: 0 = what the backend picks for bs x Hkv one-tile items (ops.extend_pick_splits), > 1 = that many.
    `page` > 1: the pool is page-granular (random page order, slots consecutive inside a page -- what SGLang's
    paged allocator hands out); P and E must be multiples of it."""
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    assert P % page == 0 and E % page == 0
    rows = (P if shared else bs * P) + bs * E + page
    kp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    vp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    if pool_dtype in (torch.float8_e4m3fn, torch.float8_e5m2):
        kp, vp = kp.to(pool_dtype).view(torch.uint8), vp.to(pool_dtype).view(torch.uint8)
    if page == 1:
        perm = torch.randperm(rows - 1, generator=g, device=dev).int() + 1
    else:
        pages = torch.randperm(rows // page - 1, generator=g, device=dev).int() + 1
        perm = (pages.view(-1, 1) * page + torch.arange(page, device=dev, dtype=torch.int32).view(1, -1)).reshape(-1)
    rtt = torch.zeros(bs + 1, P + E, dtype=torch.int32, device=dev)
    if shared:
        rtt[1:, :P] = perm[:P]
        rtt[1:, P:] = perm[P : P + bs * E].view(bs, E)
    else:
        rtt[1:] = perm[: bs * (P + E)].view(bs, P + E)
    T = bs * E
    q = torch.randn(T, Hq, D, generator=g, device=dev, dtype=torch.bfloat16)
    o = torch.empty_like(q)
    rpi = torch.arange(1, bs + 1, device=dev)
    pre = torch.full((bs,), P, dtype=torch.int32, device=dev)
    ext = torch.full((bs,), E, dtype=torch.int32, device=dev)
    sm, _, _ = ops.device_info()
    bq = ops.extend_block_q(D, D, Hq // Hkv, torch.bfloat16, pool_dtype, kernel)
    if one_tile:        # what the backend plans when no request extends by more than 128 / group tokens
        bq = ops.extend_block_q_one_tile(D, D, Hq // Hkv, torch.bfloat16, pool_dtype, kernel)
    ws = ops.ExtendWorkspace.create(bs, T // bq + bs + 1, dev, num_ctas=sm)
    ops.extend_plan(ws, ext, bq, page_size=page, req_to_token=rtt, req_pool_indices=rpi, extend_prefix_lens=pre)

    mkw = {}
    if tree_mask:   # TARGET_VERIFY shape: a chain "tree" (lower-triangular) over the E draft tokens, prefix unmasked
        m = torch.ones(bs, E, P + E, dtype=torch.uint8, device=dev)
        m[:, :, P:] = torch.tril(torch.ones(E, E, dtype=torch.uint8, device=dev))
        mkw = dict(custom_mask=m.reshape(-1), mask_skip_prefix=True,
                   mask_indptr=torch.arange(bs + 1, device=dev, dtype=torch.int64) * (E * (P + E)))

    hint = 0
    if splits == 0:       # the backend's choice: one tile per request, exact count handed over (tail-only above one wave)
        splits, hint = ops.extend_pick_splits(bs * Hkv, sm), bs
    if splits > 1:
        mkw.update(num_splits=splits, split_ws=torch.empty(ops.extend_split_ws_floats(splits, T, Hq, D), device=dev),
                   num_tiles_hint=hint)

    def fn():
        ops.extend_attention(q, o, kp, vp, rtt, rpi, pre, ext, ws, D**-0.5, kernel=kernel, kv_dtype=pool_dtype,
                             page_size=page, **mkw)

    t = time_cuda(fn, iters, warmup=2)
    if t < 500e-6:      # short-extend launches run under CUDA graphs in the engine (speculative decoding)
        t = min(t, time_cuda_graph(fn))
    flops = 4 * Hq * D * bs * (E * P + E * (E + 1) / 2)
    tf = flops / t / 1e12
    return dict(case=name, kind="extend", ms=round(t * 1e3, 3), tflops=round(tf, 1), frac_measured=round(tf / TF, 3),
                frac_nominal=round(tf / 2250, 3), page_size=page, page_box_rows=int(ws.plan[4]), splits=splits)


def store_case(name, T, Hkv, D, pool_dtype, iters):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    rows = max(T * 2, 1 << 18) + 1
    k = torch.randn(T, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    v = torch.randn(T, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    es = 1 if pool_dtype == torch.float8_e4m3fn else 2
    kc = torch.zeros(rows, Hkv, D, device=dev, dtype=torch.uint8 if es == 1 else torch.bfloat16)
    vc = torch.zeros_like(kc)
    loc = torch.randperm(rows - 1, generator=g, device=dev)[:T] + 1
    fn = lambda: ops.store_kv(k, v, kc, vc, loc, cache_dtype=pool_dtype)  # noqa: E731
    t = time_cuda(fn, iters)
    if t < 80e-6:
        t = time_cuda_graph(fn)
    nbytes = T * Hkv * 2 * D * (2 + es) + T * 8
    gbs = nbytes / t / 1e9
    return dict(case=name, kind="store", us=round(t * 1e6, 2), gbs=round(gbs, 1), frac_measured=round(gbs / HBM, 3))


def rope_store_case(name, T, Hq, Hkv, D, pool_dtype, iters):
    """Fused rotary + KV store vs the bytes it must move: q r/w, k r/w, v read, pool write."""
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    rows = max(T * 2, 1 << 18) + 1
    q = torch.randn(T, Hq * D, generator=g, device=dev, dtype=torch.bfloat16)
    k = torch.randn(T, Hkv * D, generator=g, device=dev, dtype=torch.bfloat16)
    v = torch.randn(T, Hkv * D, generator=g, device=dev, dtype=torch.bfloat16)
    es = 1 if pool_dtype == torch.float8_e4m3fn else 2
    kc = torch.zeros(rows, Hkv, D, device=dev, dtype=torch.uint8 if es == 1 else torch.bfloat16)
    vc = torch.zeros_like(kc)
    loc = torch.randperm(rows - 1, generator=g, device=dev)[:T] + 1
    pos = torch.randint(0, 32768, (T,), generator=g, device=dev)
    inv = 1.0 / (500000.0 ** (torch.arange(0, D, 2, device=dev).float() / D))
    fr = torch.arange(32768, device=dev).float()[:, None] * inv[None]
    cache = torch.cat((fr.cos(), fr.sin()), dim=-1).contiguous()
    fn = lambda: ops.rope_store_kv(pos, q, k, v, cache, D, k_cache=kc, v_cache=vc, loc=loc, cache_dtype=pool_dtype)  # noqa: E731
    t = time_cuda(fn, iters)
    if t < 80e-6:
        t = time_cuda_graph(fn)
    nbytes = T * (Hq * D * 4 + Hkv * D * (4 + es) + Hkv * D * (2 + es) + 16)
    gbs = nbytes / t / 1e9
    return dict(case=name, kind="rope+store", us=round(t * 1e6, 2), gbs=round(gbs, 1), frac_measured=round(gbs / HBM, 3))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--json", default="")
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    bf, e4 = torch.bfloat16, torch.float8_e4m3fn
    out = []

    def emit(r):
        out.append(r)
        print(json.dumps(r), flush=True)

    if a.only in ("", "decode"):
        emit(decode_case("cfg2 8B bf16 bs64 ctx4096 page1", 64, 4096, 32, 8, 128, bf, 1, a.iters))
        emit(decode_case("cfg2 8B bf16 bs64 ctx4096 page16", 64, 4096, 32, 8, 128, bf, 16, a.iters))
        emit(decode_case("cfg2 ragged U[2048,4096]", 64, 4096, 32, 8, 128, bf, 1, a.iters, ragged=True))
        emit(decode_case("cfg3 8B fp8 bs256 ctx8192", 256, 8192, 32, 8, 128, e4, 1, a.iters))
        for tp, (hq, hkv) in ((2, (32, 4)), (4, (16, 2)), (8, (8, 1))):
            for bs in (8, 32):
                emit(decode_case(f"cfg4 70B tp{tp} bs{bs} ctx32768", bs, 32768, hq, hkv, 128, bf, 1, a.iters))
        emit(decode_case("cfg4 70B tp8 fp8 bs32 ctx32768", 32, 32768, 8, 1, 128, e4, 1, a.iters))
        emit(decode_case("small bs1 ctx4096 8B", 1, 4096, 32, 8, 128, bf, 1, a.iters))
    if a.only in ("", "decode", "d64"):
        # D = 64 heads (gpt-oss 64/8, Llama-3.2-1B 32/8): bf16 pool, FP8 pool on the head-group kernel, and the
        # CUDA-core kernel FP8 pools of this head dim took before
        emit(decode_case("D64 gpt-oss heads 64/8 bf16 bs64 ctx8192", 64, 8192, 64, 8, 64, bf, 1, a.iters))
        emit(decode_case("D64 gpt-oss heads 64/8 fp8 bs64 ctx8192", 64, 8192, 64, 8, 64, e4, 1, a.iters))
        emit(decode_case("D64 gpt-oss heads 64/8 fp8 bs64 ctx8192, generic kernel", 64, 8192, 64, 8, 64, e4, 1,
                         max(3, a.iters // 10), kernel=1))
        emit(decode_case("D64 llama-3.2-1b heads 32/8 fp8 bs128 ctx8192 page16", 128, 8192, 32, 8, 64, e4, 16, a.iters))
    if a.only in ("", "extend", "verify"):
        # speculative verify (tree mask over 8 draft tokens) and short extends at several batch sizes: unsplit
        # one-tile items against the split-KV the backend picks from bs x Hkv
        it = max(5, a.iters // 3)
        for bs, P in ((64, 4096), (40, 4096), (24, 4096), (16, 4096), (4, 4096), (1, 4096), (1, 32768), (8, 16384),
                      (64, 512)):
            for sp in (1, 0):
                emit(extend_case(f"verify {bs} req x ({P} + 8 draft, tree mask), one-tile, splits {'auto' if sp == 0 else sp}",
                                 bs, P, 8, 32, 8, 128, it, shared=False, tree_mask=True, one_tile=True, splits=sp))
        for sp in (1, 0):
            emit(extend_case(f"short extend 8 req x (8192 + 16 new), one-tile, splits {'auto' if sp == 0 else sp}", 8, 8192,
                             16, 32, 8, 128, it, shared=False, one_tile=True, splits=sp))
    if a.only in ("", "store"):
        emit(store_case("decode store 64 rows 8B bf16", 64, 8, 128, bf, a.iters))
        emit(store_case("decode store 256 rows 8B fp8", 256, 8, 128, e4, a.iters))
        emit(store_case("cfg5 store 131072 rows 8B bf16", 131072, 8, 128, bf, a.iters))
        emit(store_case("cfg5 store 131072 rows 8B fp8", 131072, 8, 128, e4, a.iters))
    if a.only in ("", "store", "rope"):
        emit(rope_store_case("decode rope+store 64 rows 8B bf16", 64, 32, 8, 128, bf, a.iters))
        emit(rope_store_case("cfg5 rope+store 131072 rows 8B bf16", 131072, 32, 8, 128, bf, a.iters))
        emit(rope_store_case("cfg5 rope+store 131072 rows 8B fp8", 131072, 32, 8, 128, e4, a.iters))
    if a.only == "d256":
        it = max(3, a.iters // 10)
        emit(extend_case("D256 gemma-2 heads (16/8) 8 req x (16K shared prefix + 1K), tcgen05 one-tile", 8, 16384, 1024, 16, 8,
                         256, it))
        emit(extend_case("D256 gemma-2 heads (16/8) 8 req x (16K shared prefix + 1K), FP8-e4m3 pool", 8, 16384, 1024, 16, 8,
                         256, it, pool_dtype=e4))
        emit(extend_case("D256 gemma-2 heads (16/8) 8 req x (16K shared prefix + 1K), FP8-e5m2 pool", 8, 16384, 1024, 16, 8,
                         256, it, pool_dtype=torch.float8_e5m2))
    if a.only in ("", "extend"):
        it = max(3, a.iters // 10)
        emit(extend_case("cfg5/16 8 req x (16K shared prefix + 1K)", 8, 16384, 1024, 32, 8, 128, it))
        emit(extend_case("cfg5/16 with FP8-e4m3 KV pool", 8, 16384, 1024, 32, 8, 128, it, pool_dtype=e4))
        emit(extend_case("cfg5/16 with FP8-e5m2 KV pool", 8, 16384, 1024, 32, 8, 128, it, pool_dtype=torch.float8_e5m2))
        emit(extend_case("prefill 8 req x 4K no prefix", 8, 0, 4096, 32, 8, 128, it))
        emit(extend_case("70B tp8 4 req x (8K + 2K)", 4, 8192, 2048, 8, 1, 128, it))
        emit(extend_case("verify 64 req x (4K + 8 draft, tree mask), tcgen05 two-tile items", 64, 4096, 8, 32, 8, 128, it,
                         shared=False, tree_mask=True))
        emit(extend_case("verify 64 req x (4K + 8 draft, tree mask), tcgen05 one-tile items (backend's choice)", 64,
                         4096, 8, 32, 8, 128, it, shared=False, tree_mask=True, one_tile=True))
        emit(extend_case("short extend 64 req x (4K + 16 new), one-tile items", 64, 4096, 16, 32, 8, 128, it,
                         shared=False, one_tile=True))
        emit(extend_case("short extend 64 req x (4K + 16 new), two-tile items", 64, 4096, 16, 32, 8, 128, it,
                         shared=False))
        emit(extend_case("cfg5/16 with one-tile items (for reference)", 8, 16384, 1024, 32, 8, 128, it, one_tile=True))
        emit(extend_case("verify 64 req x (4K + 8 draft, tree mask), generic", 64, 4096, 8, 32, 8, 128, 3,
                         shared=False, tree_mask=True, kernel=1))
        emit(extend_case("D256 gemma-2 heads (16/8) 8 req x (16K shared prefix + 1K), tcgen05 one-tile", 8, 16384, 1024, 16, 8,
                         256, it))
        emit(extend_case("D256 gemma-2 heads (16/8) 8 req x (16K shared prefix + 1K), FP8-e4m3 pool", 8, 16384, 1024, 16, 8,
                         256, it, pool_dtype=e4))
        emit(extend_case("D256 gemma-2 heads (16/8), generic kernel (CUDA cores), 1 req x (4K + 256)", 1, 4096, 256, 16, 8,
                         256, 3, kernel=1))
        emit(extend_case("D64 qwen2.5-0.5b heads 8 req x (16K shared prefix + 1K)", 8, 16384, 1024, 14, 2, 64, it))
        emit(extend_case("D64 32/8 heads 8 req x (16K shared prefix + 1K)", 8, 16384, 1024, 32, 8, 64, it))
        emit(extend_case("D64 32/8 heads, generic kernel (CUDA cores), 2 req", 2, 16384, 1024, 32, 8, 64, 3, kernel=1))
        if not a.quick:     # last: these 30 ms launches run into the power cap and leave the clocks low for a while
            emit(extend_case("cfg5 no sharing 16 req x (16K + 1K)", 16, 16384, 1024, 32, 8, 128, it, shared=False))
            emit(extend_case("cfg5 128 req x (16K shared prefix + 1K)", 128, 16384, 1024, 32, 8, 128, it))
            emit(extend_case("cfg5 128 req x (16K shared prefix + 1K), page 64", 128, 16384, 1024, 32, 8, 128, it, page=64))
    if a.json:
        Path(a.json).write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
