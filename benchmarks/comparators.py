#!/usr/bin/env python
"""GPU comparators: the same synthetic inputs through this repo's kernels and through the
library kernels the reference's production backends call (flashinfer, `flashinfer_backend.py` /
`trtllm_mha_backend.py:1117,1315`).  The reference tree itself cannot travel to the GPU box, its
backends are thin wrappers around these library calls.

    python benchmarks/comparators.py [--json out.json] [--only decode|extend]

Each comparator runs in its own try/except: a library that cannot JIT-compile or lacks a cubin on the
box is recorded as {"error": ...}, never silently dropped.  Outputs are also compared (max-abs) with
ours so that the two timings are known to be of the same computation.
"""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

import torch

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "benchmarks"))
from sglang_b200 import ops  # noqa: E402
from sweep import time_cuda, time_cuda_graph  # noqa: E402


def _decode_inputs(bs, L, Hq, Hkv, D, page, fp8):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    ppr = (L + page - 1) // page
    n_pages = bs * ppr
    rows = (n_pages + 1) * page
    kp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    vp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    if fp8:
        kp, vp = kp.to(torch.float8_e4m3fn), vp.to(torch.float8_e4m3fn)
    perm = (torch.randperm(n_pages, generator=g, device=dev) + 1).view(bs, ppr)
    rtt = torch.zeros(bs + 1, ppr * page, dtype=torch.int32, device=dev)
    rtt[1:] = (perm.view(bs, ppr, 1) * page + torch.arange(page, device=dev).view(1, 1, page)).view(bs, -1).int()
    q = torch.randn(bs, Hq, D, generator=g, device=dev, dtype=torch.bfloat16)
    return q, kp, vp, rtt, perm.int()


def decode_compare(name, bs, L, Hq, Hkv, D, page, fp8, iters=30):
    q, kp, vp, rtt, pages = _decode_inputs(bs, L, Hq, Hkv, D, page, fp8)
    dev = q.device
    es = 1 if fp8 else 2
    nbytes = bs * L * Hkv * 2 * D * es + bs * L * 4 + bs * Hq * 2 * D * 2
    out = dict(case=name, kind="decode", algorithmic_bytes=nbytes)
    pool_dtype = torch.float8_e4m3fn if fp8 else torch.bfloat16

    # ---- this repo ----
    lens = torch.full((bs,), L, dtype=torch.int32, device=dev)
    rpi = torch.arange(1, bs + 1, device=dev)
    sm, _, _ = ops.device_info()
    ws = ops.DecodeWorkspace.create(bs, Hq, D, dev, num_ctas=sm * ops.decode_ctas_per_sm(D, pool_dtype))
    ops.decode_plan(ws, lens, Hkv, items_per_chunk=ops.decode_items_per_chunk(Hq, Hkv, D, D, torch.bfloat16, pool_dtype))
    o = torch.empty_like(q)
    kraw, vraw = (kp.view(torch.uint8), vp.view(torch.uint8)) if fp8 else (kp, vp)

    def ours():
        ops.decode_attention(q, o, kraw, vraw, rtt, rpi, lens, ws, D**-0.5, kv_dtype=pool_dtype)

    t = time_cuda(ours, iters)
    if t < 80e-6:
        t = time_cuda_graph(ours)
    out["b200attn"] = dict(us=round(t * 1e6, 1), GBps=round(nbytes / t / 1e9, 1))
    o_ours = o.clone()

    # ---- flashinfer BatchDecodeWithPagedKVCacheWrapper (what `--attention-backend flashinfer` runs) ----
    try:
        import flashinfer

        t0 = time.time()
        wbuf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        ppr = pages.shape[1]
        indptr = torch.arange(0, bs + 1, device=dev, dtype=torch.int32) * ppr
        indices = pages.reshape(-1).contiguous()
        last = torch.full((bs,), L - (ppr - 1) * page, dtype=torch.int32, device=dev)
        kc = kp.view(-1, page, Hkv, D)
        vc = vp.view(-1, page, Hkv, D)
        for tc in (True, False):
            key = "flashinfer_decode_tc" if tc else "flashinfer_decode_cc"
            try:
                w = flashinfer.BatchDecodeWithPagedKVCacheWrapper(wbuf, "NHD", use_tensor_cores=tc)
                w.plan(indptr, indices, last, Hq, Hkv, D, page, q_data_type=torch.bfloat16,
                       kv_data_type=pool_dtype, sm_scale=D**-0.5)
                res = w.run(q, (kc, vc))
                torch.cuda.synchronize()
                tt = time_cuda(lambda: w.run(q, (kc, vc)), iters)
                out[key] = dict(us=round(tt * 1e6, 1), GBps=round(nbytes / tt / 1e9, 1),
                                max_abs_vs_ours=round(float((res.float() - o_ours.float()).abs().max()), 5),
                                setup_s=round(time.time() - t0, 1))
            except Exception as e:  # noqa: BLE001
                out[key] = dict(error=f"{type(e).__name__}: {str(e)[:300]}")
    except Exception as e:  # noqa: BLE001
        out["flashinfer_decode"] = dict(error=f"{type(e).__name__}: {str(e)[:300]}")

    # ---- flashinfer trtllm-gen decode (what `--attention-backend trtllm_mha` runs; page_size >= 16) ----
    if page >= 16 and not fp8:
        try:
            import flashinfer

            t0 = time.time()
            wbuf2 = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
            # HND view of the same NHD pool, as trtllm_mha_backend.py:1008-1025 builds it (no copy)
            kv_k = kp.view(-1, page, Hkv, D).permute(0, 2, 1, 3)
            kv_v = vp.view(-1, page, Hkv, D).permute(0, 2, 1, 3)
            bt = pages.contiguous()
            sl = torch.full((bs,), L, dtype=torch.int32, device=dev)

            def trt():
                return flashinfer.decode.trtllm_batch_decode_with_kv_cache(
                    query=q, kv_cache=(kv_k, kv_v), workspace_buffer=wbuf2, block_tables=bt, seq_lens=sl,
                    max_seq_len=L, bmm1_scale=D**-0.5, bmm2_scale=1.0)

            res = trt()
            torch.cuda.synchronize()
            tt = time_cuda(trt, iters)
            out["flashinfer_trtllm_gen_decode"] = dict(
                us=round(tt * 1e6, 1), GBps=round(nbytes / tt / 1e9, 1),
                max_abs_vs_ours=round(float((res.float().view_as(o_ours) - o_ours.float()).abs().max()), 5),
                setup_s=round(time.time() - t0, 1))
        except Exception as e:  # noqa: BLE001
            out["flashinfer_trtllm_gen_decode"] = dict(error=f"{type(e).__name__}: {str(e)[:300]}")
    return out


def extend_compare(name, bs, P, E, Hq, Hkv, D, iters=5):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    rows = P + bs * E + 1
    kp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    vp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    perm = torch.randperm(rows - 1, generator=g, device=dev).int() + 1
    rtt = torch.zeros(bs + 1, P + E, dtype=torch.int32, device=dev)
    rtt[1:, :P] = perm[:P]
    rtt[1:, P:] = perm[P: P + bs * E].view(bs, E)
    T = bs * E
    q = torch.randn(T, Hq, D, generator=g, device=dev, dtype=torch.bfloat16)
    flops = 4 * Hq * D * bs * (E * P + E * (E + 1) / 2)
    out = dict(case=name, kind="extend", algorithmic_flops=flops)

    o = torch.empty_like(q)
    rpi = torch.arange(1, bs + 1, device=dev)
    pre = torch.full((bs,), P, dtype=torch.int32, device=dev)
    ext = torch.full((bs,), E, dtype=torch.int32, device=dev)
    sm, _, _ = ops.device_info()
    bq = ops.extend_block_q(D, D, Hq // Hkv, torch.bfloat16, torch.bfloat16, 0)
    ws = ops.ExtendWorkspace.create(bs, T // bq + bs + 1, dev, num_ctas=sm)
    ops.extend_plan(ws, ext, bq)

    def ours():
        ops.extend_attention(q, o, kp, vp, rtt, rpi, pre, ext, ws, D**-0.5, kernel=0, kv_dtype=torch.bfloat16)

    t = time_cuda(ours, iters, warmup=2)
    out["b200attn"] = dict(ms=round(t * 1e3, 3), TFLOPs=round(flops / t / 1e12, 1))
    o_ours = o.clone()

    try:
        import flashinfer

        wbuf = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
        qo_indptr = torch.arange(0, bs + 1, device=dev, dtype=torch.int32) * E
        kv_indptr = torch.arange(0, bs + 1, device=dev, dtype=torch.int32) * (P + E)
        kv_indices = rtt[1:].reshape(-1).contiguous()
        last = torch.ones(bs, dtype=torch.int32, device=dev)
        kc, vc = kp.view(-1, 1, Hkv, D), vp.view(-1, 1, Hkv, D)
        for backend in ("auto", "fa2"):
            key = f"flashinfer_prefill_{backend}"
            try:
                t0 = time.time()
                w = flashinfer.BatchPrefillWithPagedKVCacheWrapper(wbuf, "NHD", backend=backend)
                w.plan(qo_indptr, kv_indptr, kv_indices, last, Hq, Hkv, D, 1, causal=True, sm_scale=D**-0.5,
                       q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
                res = w.run(q, (kc, vc))
                torch.cuda.synchronize()
                tt = time_cuda(lambda: w.run(q, (kc, vc)), iters, warmup=1)
                out[key] = dict(ms=round(tt * 1e3, 3), TFLOPs=round(flops / tt / 1e12, 1),
                                max_abs_vs_ours=round(float((res.float() - o_ours.float()).abs().max()), 5),
                                setup_s=round(time.time() - t0, 1))
            except Exception as e:  # noqa: BLE001
                out[key] = dict(error=f"{type(e).__name__}: {str(e)[:300]}")
    except Exception as e:  # noqa: BLE001
        out["flashinfer_prefill"] = dict(error=f"{type(e).__name__}: {str(e)[:300]}")
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=None)
    ap.add_argument("--only", default=None)
    args = ap.parse_args()
    res = []

    def emit(r):
        res.append(r)
        print(json.dumps(r), flush=True)

    if args.only in (None, "decode"):
        emit(decode_compare("cfg2 8B bf16 bs64 ctx4096 page1", 64, 4096, 32, 8, 128, 1, False))
        emit(decode_compare("cfg2 8B bf16 bs64 ctx4096 page16", 64, 4096, 32, 8, 128, 16, False))
        emit(decode_compare("cfg3 8B fp8 bs256 ctx8192 page1", 256, 8192, 32, 8, 128, 1, True, iters=10))
        emit(decode_compare("cfg4 70B tp8 bs32 ctx32768 page1", 32, 32768, 8, 1, 128, 1, False))
    if args.only in (None, "extend"):
        emit(extend_compare("cfg5/16 8 req x (16K shared prefix + 1K)", 8, 16384, 1024, 32, 8, 128))
    if args.json:
        Path(args.json).write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
