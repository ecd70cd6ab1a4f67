#!/usr/bin/env python
"""GPU comparators: the same synthetic inputs through this repo's kernels and through the
library kernels the reference's production backends call (flashinfer, `flashinfer_backend.py` /
`trtllm_mha_backend.py:1117,1315`), and -- through the staged reference tree `oracle/_ref` -- the reference's own
Triton kernels (K6-K9: `kernels/ops/attention/decode_attention.py:968-1044`, `extend_attention.py:664-812`) and its
vendored FA4 (K12: `kernels/ops/attention/flash_attn/cute/interface.py`, paged KV through `paged_kv.py`).

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


def _ref_on_path():
    ref = REPO / "oracle" / "_ref"
    if not (ref / "MANIFEST.json").exists():
        raise RuntimeError("oracle/_ref is not staged")
    for p_ in (str(ref / "stubs"), str(ref / "python")):
        if p_ not in sys.path:
            sys.path.insert(0, p_)


def _paged_extend_inputs(bs, P, E, Hq, Hkv, D, page):
    """Shared-prefix extend problem on a page-granular pool (prefix pages identical in every request)."""
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    assert P % page == 0 and E % page == 0
    n_pages = P // page + bs * (E // page)
    rows = (n_pages + 1) * page
    kp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    vp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    perm = (torch.randperm(n_pages, generator=g, device=dev) + 1).int()
    pages = torch.empty(bs, (P + E) // page, dtype=torch.int32, device=dev)
    pages[:, : P // page] = perm[: P // page]
    pages[:, P // page :] = perm[P // page :].view(bs, E // page)
    rtt = torch.zeros(bs + 1, P + E, dtype=torch.int32, device=dev)
    rtt[1:] = (pages.view(bs, -1, 1) * page + torch.arange(page, device=dev, dtype=torch.int32).view(1, 1, page)).view(bs, -1)
    q = torch.randn(bs * E, Hq, D, generator=g, device=dev, dtype=torch.bfloat16)
    return q, kp, vp, rtt, pages


def extend_compare_blackwell(name, bs, P, E, Hq, Hkv, D, page, iters=5, ours_only=False):
    """Ours vs the shipped Blackwell kernels on the same page-granular pool: trtllm-gen context attention
    (flashinfer.prefill.trtllm_batch_context_with_kv_cache, trtllm_mha_backend.py:1315), the reference's vendored FA4
    (flash_attn/cute/interface.py flash_attn_varlen_func with page_table) and the reference's Triton extend kernel."""
    dev = "cuda"
    q, kp, vp, rtt, pages = _paged_extend_inputs(bs, P, E, Hq, Hkv, D, page)
    T = bs * E
    flops = 4 * Hq * D * bs * (E * P + E * (E + 1) / 2)
    out = dict(case=name, kind="extend", page_size=page, algorithmic_flops=flops)
    o = torch.empty_like(q)
    rpi = torch.arange(1, bs + 1, device=dev)
    pre = torch.full((bs,), P, dtype=torch.int32, device=dev)
    ext = torch.full((bs,), E, dtype=torch.int32, device=dev)
    sm, _, _ = ops.device_info()
    bq = ops.extend_block_q(D, D, Hq // Hkv, torch.bfloat16, torch.bfloat16, 0)
    ws = ops.ExtendWorkspace.create(bs, T // bq + bs + 1, dev, num_ctas=sm)
    ops.extend_plan(ws, ext, bq, page_size=page, req_to_token=rtt, req_pool_indices=rpi, extend_prefix_lens=pre)

    def ours(pg=page):
        ops.extend_attention(q, o, kp, vp, rtt, rpi, pre, ext, ws, D**-0.5, kernel=0, kv_dtype=torch.bfloat16,
                             page_size=pg)

    # what the backend runs at this page size (TMA page boxes for page >= 16, LDGSTS row gather for page 1) ...
    t = time_cuda(ours, iters, warmup=2)
    out["b200attn"] = dict(ms=round(t * 1e3, 3), TFLOPs=round(flops / t / 1e12, 1),
                           fill="tma_page_boxes" if page >= 16 else "ldgsts_row_gather")
    o_ours = o.clone()
    if page >= 16:  # ... and the row gather on the same pool, for the A/B
        t1 = time_cuda(lambda: ours(1), iters, warmup=2)
        out["b200attn_row_gather"] = dict(ms=round(t1 * 1e3, 3), TFLOPs=round(flops / t1 / 1e12, 1),
                                          max_abs_vs_page_fill=round(float((o.float() - o_ours.float()).abs().max()), 6))
    if ours_only:
        return out
    cu_q = torch.arange(0, bs + 1, device=dev, dtype=torch.int32) * E
    cu_k = torch.arange(0, bs + 1, device=dev, dtype=torch.int32) * (P + E)
    seq_k = torch.full((bs,), P + E, dtype=torch.int32, device=dev)

    def record(key, fn, t0):
        try:
            res = fn()
            torch.cuda.synchronize()
            tt = time_cuda(fn, iters, warmup=1)
            res = res[0] if isinstance(res, tuple) else res
            out[key] = dict(ms=round(tt * 1e3, 3), TFLOPs=round(flops / tt / 1e12, 1),
                            max_abs_vs_ours=round(float((res.float().view_as(o_ours) - o_ours.float()).abs().max()), 5),
                            setup_s=round(time.time() - t0, 1))
        except Exception as e:  # noqa: BLE001
            out[key] = dict(error=f"{type(e).__name__}: {str(e)[:400]}")

    # ---- trtllm-gen context attention (page_size >= 16) ----
    if page >= 16:
        try:
            import flashinfer

            t0 = time.time()
            wbuf = torch.zeros(512 << 20, dtype=torch.uint8, device=dev)
            kv_k = kp.view(-1, page, Hkv, D).permute(0, 2, 1, 3)       # HND view of the NHD pool, no copy
            kv_v = vp.view(-1, page, Hkv, D).permute(0, 2, 1, 3)       # (trtllm_mha_backend.py:1008-1025)
            bt = pages.contiguous()
            record("flashinfer_trtllm_gen_context", lambda: flashinfer.prefill.trtllm_batch_context_with_kv_cache(
                query=q, kv_cache=(kv_k, kv_v), workspace_buffer=wbuf, block_tables=bt, seq_lens=seq_k, max_q_len=E,
                max_kv_len=P + E, bmm1_scale=D**-0.5, bmm2_scale=1.0, batch_size=bs, cum_seq_lens_q=cu_q,
                cum_seq_lens_kv=cu_k, out_dtype=torch.bfloat16), t0)
        except Exception as e:  # noqa: BLE001
            out["flashinfer_trtllm_gen_context"] = dict(error=f"{type(e).__name__}: {str(e)[:400]}")

    # ---- the reference's vendored FA4 (CuTe DSL) with a page table ----
    try:
        _ref_on_path()
        t0 = time.time()
        from sglang.kernels.ops.attention.flash_attn.cute.interface import flash_attn_varlen_func

        k4, v4 = kp.view(-1, page, Hkv, D), vp.view(-1, page, Hkv, D)
        record("reference_fa4_cute_paged", lambda: flash_attn_varlen_func(
            q, k4, v4, cu_seqlens_q=cu_q, seqused_k=seq_k, max_seqlen_q=E, max_seqlen_k=P + E,
            page_table=pages.contiguous(), softmax_scale=D**-0.5, causal=True), t0)
    except Exception as e:  # noqa: BLE001
        out["reference_fa4_cute_paged"] = dict(error=f"{type(e).__name__}: {str(e)[:400]}")

    # ---- the reference's Triton extend kernel (what --attention-backend triton runs) ----
    try:
        _ref_on_path()
        t0 = time.time()
        from sglang.kernels.ops.attention.extend_attention import extend_attention_fwd

        kv_indptr = torch.arange(0, bs + 1, device=dev, dtype=torch.int32) * P       # prefix only (triton_backend.py:876-892)
        kv_indices = rtt[1:, :P].reshape(-1).to(torch.int64).contiguous()
        new_loc = rtt[1:, P:].reshape(-1).long()
        k_ext, v_ext = kp[new_loc].contiguous(), vp[new_loc].contiguous()
        o_t = torch.empty_like(q)

        def tri():
            extend_attention_fwd(q, k_ext, v_ext, o_t, kp, vp, cu_q, kv_indptr, kv_indices, None, True, None, E,
                                 1.0, 1.0, D**-0.5)
            return o_t

        record("reference_triton_extend", tri, t0)
    except Exception as e:  # noqa: BLE001
        out["reference_triton_extend"] = dict(error=f"{type(e).__name__}: {str(e)[:400]}")
    return out


def decode_compare_triton(name, bs, L, Hq, Hkv, D, iters=20):
    """Ours vs the reference's Triton split-KV decode (decode_attention.py:968-1044) on a page-1 pool, with the
    metadata the Triton backend builds (kv_indptr / kv_indices copy, num_kv_splits; triton_backend.py:304-351)."""
    q, kp, vp, rtt, _ = _decode_inputs(bs, L, Hq, Hkv, D, 1, False)
    dev = q.device
    nbytes = bs * L * Hkv * 2 * D * 2 + bs * L * 4 + bs * Hq * 2 * D * 2
    out = dict(case=name, kind="decode", algorithmic_bytes=nbytes)
    lens = torch.full((bs,), L, dtype=torch.int32, device=dev)
    rpi = torch.arange(1, bs + 1, device=dev)
    sm, _, _ = ops.device_info()
    ws = ops.DecodeWorkspace.create(bs, Hq, D, dev, num_ctas=sm * ops.decode_ctas_per_sm(D, torch.bfloat16))
    ops.decode_plan(ws, lens, Hkv, items_per_chunk=ops.decode_items_per_chunk(Hq, Hkv, D, D, torch.bfloat16, torch.bfloat16))
    o = torch.empty_like(q)
    ours = lambda: ops.decode_attention(q, o, kp, vp, rtt, rpi, lens, ws, D**-0.5)   # noqa: E731
    t_e = time_cuda(ours, iters)
    t_g = time_cuda_graph(ours)
    out["b200attn"] = dict(us_eager=round(t_e * 1e6, 1), us_replay=round(t_g * 1e6, 1), GBps_replay=round(nbytes / t_g / 1e9, 1))
    o_ours = o.clone()
    try:
        _ref_on_path()
        t0 = time.time()
        from sglang.kernels.ops.attention.decode_attention import decode_attention_fwd

        kv_indptr = torch.arange(0, bs + 1, device=dev, dtype=torch.int32) * L
        kv_indices = rtt[1:].reshape(-1).to(torch.int64).contiguous()
        for splits in (8, 16, 32):
            attn_logits = torch.empty(bs, Hq, splits, D, dtype=torch.float32, device=dev)
            attn_lse = torch.empty(bs, Hq, splits, dtype=torch.float32, device=dev)
            nks = torch.full((bs,), splits, dtype=torch.int32, device=dev)
            o_t = torch.empty_like(q)
            tri = lambda: decode_attention_fwd(q, kp, vp, o_t, kv_indptr, kv_indices, attn_logits, attn_lse, nks,  # noqa: E731
                                               splits, D**-0.5, 1.0, 1.0)
            try:
                tri()
                torch.cuda.synchronize()
                te, tg = time_cuda(tri, iters), time_cuda_graph(tri)
                out[f"reference_triton_decode_splits{splits}"] = dict(
                    us_eager=round(te * 1e6, 1), us_replay=round(tg * 1e6, 1), GBps_replay=round(nbytes / tg / 1e9, 1),
                    max_abs_vs_ours=round(float((o_t.float() - o_ours.float()).abs().max()), 5),
                    setup_s=round(time.time() - t0, 1))
            except Exception as e:  # noqa: BLE001
                out[f"reference_triton_decode_splits{splits}"] = dict(error=f"{type(e).__name__}: {str(e)[:400]}")
    except Exception as e:  # noqa: BLE001
        out["reference_triton_decode"] = dict(error=f"{type(e).__name__}: {str(e)[:400]}")
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
    if args.only in (None, "blackwell", "ours"):
        oo = args.only == "ours"
        for page in (1, 16, 64, 128):
            emit(extend_compare_blackwell(f"cfg5/16 8 req x (16K shared prefix + 1K), page {page}", 8, 16384, 1024,
                                          32, 8, 128, page, ours_only=oo))
        emit(extend_compare_blackwell("prefill 8 req x 4K no prefix, page 64", 8, 0, 4096, 32, 8, 128, 64,
                                      ours_only=oo))
        emit(extend_compare_blackwell("cfg5 128 req x (16K shared prefix + 1K), page 64", 128, 16384, 1024, 32, 8,
                                      128, 64, iters=3, ours_only=oo))
        emit(extend_compare_blackwell("D256 gemma-2 heads (16/8) 8 req x (16K shared prefix + 1K), page 64", 8, 16384,
                                      1024, 16, 8, 256, 64, ours_only=oo))
    if args.only in (None, "triton"):
        emit(decode_compare_triton("cfg2 8B bf16 bs64 ctx4096 page1", 64, 4096, 32, 8, 128))
        emit(decode_compare_triton("cfg4 70B tp8 bs32 ctx32768 page1", 32, 32768, 8, 1, 128))
    if args.json:
        Path(args.json).write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
