"""Generate tests/golden/*.npz by running the REFERENCE itself on CPU (test infrastructure).

Run in the build container only (needs /root/reference; the GPU box uses the committed files):

    PYTHONPATH=oracle/stubs:/root/reference/python python oracle/gen_golden.py

What runs is the reference's own code, unmodified, imported from /root/reference/python:
  - `MHATokenToKVPool.set_kv_buffer`, `ReqToTokenPool`          (srt/mem_cache/memory_pool.py)
  - `TorchNativeAttnBackend.forward_decode / forward_extend`     (srt/layers/attention/torch_native_backend.py)
  - the dense attention unit-test kit's 12 cases and slot layouts
    (test/kits/attention_unittest/attention_methods/dense_attention.py)
  - torch's own bf16/f16 -> float8 casts for the exhaustive cast tables
Outputs (all arrays stored as raw integer bit patterns so nothing depends on numpy dtypes):
  tests/golden/fp8_cast_tables.npz    4 x uint8[65536]
  tests/golden/kv_store.npz           set_kv_buffer inputs and resulting pool rows
  tests/golden/attention.npz          decode/extend inputs, stored pool rows, reference outputs in the model dtype,
                                      and the same reference code run in float32
"""
from __future__ import annotations

import sys
import unittest
from pathlib import Path
from types import SimpleNamespace

import numpy as np
import torch

REPO = Path(__file__).resolve().parent.parent
GOLDEN = REPO / "tests" / "golden"

from sglang.srt.layers.attention.torch_native_backend import TorchNativeAttnBackend  # noqa: E402
from sglang.srt.layers.radix_attention import RadixAttention  # noqa: E402
from sglang.srt.mem_cache.memory_pool import KVWriteLoc, MHATokenToKVPool, ReqToTokenPool  # noqa: E402
from sglang.srt.model_executor.forward_batch_info import ForwardBatch, ForwardMode  # noqa: E402
from sglang.srt.model_executor.forward_context import ForwardContext, forward_context  # noqa: E402
from sglang.test.kits.attention_unittest.attention_methods import dense_attention as kit  # noqa: E402


def bits(t: torch.Tensor) -> np.ndarray:
    """Raw bit pattern of a tensor as an unsigned integer array."""
    t = t.detach().contiguous().cpu()
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    if t.dtype in (torch.float8_e4m3fn, torch.float8_e5m2):
        return t.view(torch.uint8).numpy().copy()
    return t.numpy().copy()


# ----------------------------------------------------------------------------- cast tables
def gen_cast_tables():
    pat = torch.arange(65536, dtype=torch.int32).to(torch.int16)
    out = {}
    for sname, sdt in (("bf16", torch.bfloat16), ("f16", torch.float16)):
        src = pat.view(sdt)
        for dname, ddt in (("e4m3", torch.float8_e4m3fn), ("e5m2", torch.float8_e5m2)):
            out[f"{sname}_to_{dname}"] = src.to(ddt).view(torch.uint8).numpy().copy()
    np.savez_compressed(GOLDEN / "fp8_cast_tables.npz", **out)
    print("fp8_cast_tables.npz", {k: v.shape for k, v in out.items()})


# ----------------------------------------------------------------------------- KV store
def make_pool(size, page_size, dtype, head_num, head_dim, v_head_dim=None):
    return MHATokenToKVPool(
        size=size, page_size=page_size, dtype=dtype, head_num=head_num, head_dim=head_dim, layer_num=1,
        device="cpu", enable_memory_saver=False, enable_alt_stream=False,
        **({"v_head_dim": v_head_dim} if v_head_dim is not None else {}),
    )


def gen_kv_store():
    out = {}
    cases = [
        # name, src dtype, pool dtype, T, Hkv, D, k_scale, v_scale, value std
        ("bf16_copy", torch.bfloat16, torch.bfloat16, 37, 2, 64, None, None, 1.0),
        ("f16_copy", torch.float16, torch.float16, 19, 4, 16, None, None, 1.0),
        ("bf16_to_e4m3", torch.bfloat16, torch.float8_e4m3fn, 41, 2, 64, None, None, 1.0),
        ("bf16_to_e4m3_wide", torch.bfloat16, torch.float8_e4m3fn, 33, 1, 128, None, None, 200.0),
        ("bf16_to_e4m3_scaled_pow2", torch.bfloat16, torch.float8_e4m3fn, 29, 2, 64, 2.0, 0.5, 3.0),
        ("bf16_to_e4m3_scaled", torch.bfloat16, torch.float8_e4m3fn, 31, 2, 64, 0.37, 1.9, 1.0),
        ("f16_to_e4m3_scaled", torch.float16, torch.float8_e4m3fn, 23, 2, 32, 0.83, 1.3, 1.0),
        ("bf16_to_e5m2", torch.bfloat16, torch.float8_e5m2, 27, 2, 64, None, None, 50.0),
        ("bf16_odd_width", torch.bfloat16, torch.bfloat16, 9, 1, 13, None, None, 1.0),
    ]
    names = []
    g = torch.Generator().manual_seed(1234)
    for name, sdt, pdt, T, H, D, ks, vs, std in cases:
        size, page = 96, 1
        pool = make_pool(size, page, pdt, H, D)
        k = (torch.randn(T, H, D, generator=g) * std).to(sdt)
        v = (torch.randn(T, H, D, generator=g) * std).to(sdt)
        if "wide" in name:  # exercise the non-saturating overflow rule and tiny values
            k.view(-1)[:12] = torch.tensor([448, 449, 464, 465, 466, 479, 480, 1000, -464, -466, 1e-4, 2**-10]).to(sdt)
            v.view(-1)[:4] = torch.tensor([float("inf"), float("-inf"), float("nan"), 2**-9]).to(sdt)
        loc = torch.randperm(size + page - 1, generator=g)[:T] + 1  # never slot 0, no duplicates
        layer = SimpleNamespace(layer_id=0)
        # the reference scales in place: keep pristine copies as the recorded inputs
        k_in, v_in = k.clone(), v.clone()
        if ks is None:
            pool.set_kv_buffer(layer, KVWriteLoc(loc), k, v)
        else:
            pool.set_kv_buffer(layer, KVWriteLoc(loc), k, v, ks, vs)
        names.append(name)
        out[f"{name}.k"] = bits(k_in)
        out[f"{name}.v"] = bits(v_in)
        out[f"{name}.loc"] = loc.numpy().astype(np.int64)
        out[f"{name}.k_cache"] = bits(pool.k_buffer[0])
        out[f"{name}.v_cache"] = bits(pool.v_buffer[0])
        out[f"{name}.meta"] = np.array(
            [kit_dtype_code(sdt), kit_dtype_code(pdt), T, H, D, size + page], dtype=np.int64)
        out[f"{name}.scales"] = np.array([ks or 1.0, vs or 1.0], dtype=np.float64)
    out["names"] = np.array(names)
    np.savez_compressed(GOLDEN / "kv_store.npz", **out)
    print("kv_store.npz", names)


def kit_dtype_code(dt):
    return {torch.bfloat16: 0, torch.float16: 1, torch.float8_e4m3fn: 2, torch.float8_e5m2: 3,
            torch.float32: 4}[dt]


# ----------------------------------------------------------------------------- attention
def run_reference_attention(mode, dtype, pool_dtype, Hq, Hkv, D, page_size, req_to_token, req_pool_indices,
                            prefix_lens, extend_lens, q, k_new, v_new, out_cache_loc, k_pool_init, v_pool_init,
                            pool_rows, save_kv_cache=True):
    """Drive the reference TorchNativeAttnBackend with the reference pools on CPU."""
    bs = len(prefix_lens)
    seq_lens = [p + e for p, e in zip(prefix_lens, extend_lens)]
    rtp = ReqToTokenPool(size=req_to_token.shape[0] - 1, max_context_len=req_to_token.shape[1], device="cpu",
                         enable_memory_saver=False)
    rtp.req_to_token.copy_(req_to_token)
    pool = make_pool(pool_rows - page_size, page_size, pool_dtype, Hkv, D)
    pool.k_buffer[0].copy_(k_pool_init)
    pool.v_buffer[0].copy_(v_pool_init)
    backend = TorchNativeAttnBackend(SimpleNamespace(device="cpu", req_to_token_pool=rtp, token_to_kv_pool=pool))
    layer = RadixAttention(num_heads=Hq, head_dim=D, scaling=D**-0.5, num_kv_heads=Hkv, layer_id=0)
    fb = ForwardBatch(
        forward_mode=mode, batch_size=bs,
        input_ids=torch.arange(q.shape[0], dtype=torch.int64),
        req_pool_indices=req_pool_indices,
        seq_lens=torch.tensor(seq_lens, dtype=torch.int32),
        seq_lens_cpu=torch.tensor(seq_lens, dtype=torch.int32),
        out_cache_loc=out_cache_loc, seq_lens_sum=sum(seq_lens),
        positions=torch.zeros(q.shape[0], dtype=torch.int64),
    )
    if mode.is_extend():
        fb.extend_prefix_lens = torch.tensor(prefix_lens, dtype=torch.int32)
        fb.extend_prefix_lens_cpu = list(prefix_lens)
        fb.extend_seq_lens = torch.tensor(extend_lens, dtype=torch.int32)
        fb.extend_seq_lens_cpu = list(extend_lens)
        fb.extend_num_tokens = sum(extend_lens)
    with torch.no_grad(), forward_context(ForwardContext(attn_backend=backend)):
        backend.init_forward_metadata(fb)
        o = backend.forward(q.reshape(q.shape[0], -1).clone(), k_new.clone(), v_new.clone(), layer, fb,
                            save_kv_cache=save_kv_cache)
    return o, pool


def synth_case(name, mode, dtype, pool_dtype, Hq, Hkv, D, page_size, prefix_lens, extend_lens, shared_prefix=0,
               seed=0):
    """Synthetic BASELINE-shaped case (scaled down): random slot permutation, N(0,1) values."""
    g = torch.Generator().manual_seed(seed)
    bs = len(prefix_lens)
    seq_lens = [p + e for p, e in zip(prefix_lens, extend_lens)]
    max_ctx = max(seq_lens) + 3
    n_pages_needed = sum((L - shared_prefix + page_size - 1) // page_size for L in seq_lens) \
        + (shared_prefix + page_size - 1) // page_size + 2
    n_pages = n_pages_needed + 5
    pool_rows = (n_pages + 1) * page_size  # page 0 reserved
    perm = (torch.randperm(n_pages, generator=g) + 1).tolist()
    req_to_token = torch.zeros(bs + 1, max_ctx, dtype=torch.int32)
    req_pool_indices = torch.arange(1, bs + 1, dtype=torch.int64)
    shared_pages = [perm.pop() for _ in range((shared_prefix + page_size - 1) // page_size)] if shared_prefix else []
    for r, L in enumerate(seq_lens):
        pos = 0
        if shared_prefix:
            assert shared_prefix % page_size == 0 and shared_prefix <= prefix_lens[r]
            for pg in shared_pages:
                req_to_token[r + 1, pos : pos + page_size] = torch.arange(pg * page_size, (pg + 1) * page_size)
                pos += page_size
        while pos < L:
            pg = perm.pop()
            n = min(page_size, L - pos)
            req_to_token[r + 1, pos : pos + n] = torch.arange(pg * page_size, pg * page_size + n)
            pos += n
    T = sum(extend_lens)
    q = torch.randn(T, Hq, D, generator=g).to(dtype)
    k_new = torch.randn(T, Hkv, D, generator=g).to(dtype)
    v_new = torch.randn(T, Hkv, D, generator=g).to(dtype)
    out_cache_loc = torch.cat([
        req_to_token[r + 1, prefix_lens[r] : seq_lens[r]].long() for r in range(bs)])
    # prefix content: random rows everywhere, cast like the pool would
    k_pool = torch.randn(pool_rows, Hkv, D, generator=g).to(dtype)
    v_pool = torch.randn(pool_rows, Hkv, D, generator=g).to(dtype)
    if pool_dtype in (torch.float8_e4m3fn, torch.float8_e5m2):
        k_pool_store = k_pool.to(pool_dtype).view(torch.uint8)
        v_pool_store = v_pool.to(pool_dtype).view(torch.uint8)
    else:
        k_pool_store, v_pool_store = k_pool.to(pool_dtype), v_pool.to(pool_dtype)
    return dict(name=name, mode=mode, dtype=dtype, pool_dtype=pool_dtype, Hq=Hq, Hkv=Hkv, D=D, page_size=page_size,
                prefix_lens=list(prefix_lens), extend_lens=list(extend_lens), req_to_token=req_to_token,
                req_pool_indices=req_pool_indices, q=q, k_new=k_new, v_new=v_new, out_cache_loc=out_cache_loc,
                k_pool=k_pool_store, v_pool=v_pool_store, pool_rows=pool_rows)


def kit_case(case, layout):
    """One of the reference kit's 12 dense cases, captured at the RadixAttention boundary."""
    t = unittest.TestCase()
    fx = kit.build_dense_attention_fixture(t, case, dtype=torch.bfloat16, device="cpu", loc_layout=layout)
    q, k, v = fx.actual_module.project_qkv(fx.input_hidden)
    pool, rtp, fb = fx.runner.token_to_kv_pool, fx.runner.req_to_token_pool, fx.forward_batch
    D = kit.DEFAULT_HEAD_DIM
    extend_lens = list(case.input_lens)
    return dict(name=f"kit.{case.name}.{layout}", mode=case.forward_mode, dtype=torch.bfloat16,
                pool_dtype=torch.bfloat16, Hq=case.num_heads, Hkv=case.num_kv_heads, D=D, page_size=case.page_size,
                prefix_lens=list(case.prefix_lens), extend_lens=extend_lens,
                req_to_token=rtp.req_to_token.clone(), req_pool_indices=fb.req_pool_indices.to(torch.int64),
                q=q.detach().view(-1, case.num_heads, D), k_new=k.detach().view(-1, case.num_kv_heads, D),
                v_new=v.detach().view(-1, case.num_kv_heads, D), out_cache_loc=fb.out_cache_loc.clone(),
                k_pool=pool.k_buffer[0].clone(), v_pool=pool.v_buffer[0].clone(),
                pool_rows=pool.k_buffer[0].shape[0])


def gen_attention():
    cases = []
    for c in kit.make_dense_cases("torch_native"):
        for layout in ("shuffled_pages", "interleaved_pages", "non_monotonic_extend"):
            cases.append(kit_case(c, layout))
    DEC, EXT = ForwardMode.DECODE, ForwardMode.EXTEND
    bf, f16, e4 = torch.bfloat16, torch.float16, torch.float8_e4m3fn
    # BASELINE.json configs, scaled so the committed fixtures stay small
    cases += [
        # cfg1: Qwen2.5-0.5B heads, bs 8 decode, 512-token shared prefix, page 1 (full size)
        synth_case("cfg1.qwen_decode_shared_prefix", DEC, bf, bf, 14, 2, 64, 1,
                   [512 + n for n in (0, 5, 17, 31, 40, 47, 58, 63)], [1] * 8, shared_prefix=512, seed=1),
        # cfg2: Llama-8B group/head_dim (Hq/Hkv = 4, D 128), ragged, page 1 and page 16
        synth_case("cfg2.llama_decode_page1", DEC, bf, bf, 8, 2, 128, 1, [199, 64, 1, 130], [1] * 4, seed=2),
        synth_case("cfg2.llama_decode_page16", DEC, bf, bf, 8, 2, 128, 16, [175, 16, 31, 120], [1] * 4, seed=3),
        # cfg3: FP8-e4m3 pool
        synth_case("cfg3.llama_decode_fp8", DEC, bf, e4, 8, 2, 128, 1, [200, 97, 33], [1] * 3, seed=4),
        # cfg4: 70B TP8 shard: 8 q heads on 1 kv head, long context
        synth_case("cfg4.llama70b_tp8_decode", DEC, bf, bf, 8, 1, 128, 1, [420, 133], [1] * 2, seed=5),
        # cfg5: shared prefix + private extend
        synth_case("cfg5.radix_extend_shared_prefix", EXT, bf, bf, 8, 2, 128, 1, [96] * 4, [40, 33, 17, 40],
                   shared_prefix=96, seed=6),
        synth_case("cfg5.radix_extend_page16_fp8", EXT, bf, e4, 8, 2, 128, 16, [64, 32], [37, 50],
                   shared_prefix=32, seed=7),
        synth_case("f16.decode_d64", DEC, f16, f16, 4, 4, 64, 1, [77, 140], [1] * 2, seed=8),
        synth_case("f16.extend_d64_mqa", EXT, f16, f16, 4, 1, 64, 1, [0, 30], [45, 20], seed=9),
    ]
    out, names = {}, []
    for c in cases:
        is_fp8 = c["pool_dtype"] in (torch.float8_e4m3fn, torch.float8_e5m2)
        o, pool = run_reference_attention(
            c["mode"], c["dtype"], c["pool_dtype"], c["Hq"], c["Hkv"], c["D"], c["page_size"], c["req_to_token"],
            c["req_pool_indices"], c["prefix_lens"], c["extend_lens"], c["q"], c["k_new"], c["v_new"],
            c["out_cache_loc"], c["k_pool"], c["v_pool"], c["pool_rows"])
        # the same reference code in float32, on the values the pool actually holds after the store
        k_after, v_after = pool.get_key_buffer(0), pool.get_value_buffer(0)
        o32, _ = run_reference_attention(
            c["mode"], torch.float32, torch.float32, c["Hq"], c["Hkv"], c["D"], c["page_size"], c["req_to_token"],
            c["req_pool_indices"], c["prefix_lens"], c["extend_lens"], c["q"].float(), c["k_new"].float(),
            c["v_new"].float(), c["out_cache_loc"], k_after.float(), v_after.float(), c["pool_rows"],
            save_kv_cache=False)
        n = c["name"]
        names.append(n)
        out[f"{n}.meta"] = np.array([
            1 if c["mode"].is_decode() else 0, kit_dtype_code(c["dtype"]), kit_dtype_code(c["pool_dtype"]),
            c["Hq"], c["Hkv"], c["D"], c["page_size"], c["pool_rows"]], dtype=np.int64)
        out[f"{n}.prefix_lens"] = np.array(c["prefix_lens"], dtype=np.int64)
        out[f"{n}.extend_lens"] = np.array(c["extend_lens"], dtype=np.int64)
        out[f"{n}.req_to_token"] = c["req_to_token"].numpy().astype(np.int32)
        out[f"{n}.req_pool_indices"] = c["req_pool_indices"].numpy().astype(np.int64)
        out[f"{n}.out_cache_loc"] = c["out_cache_loc"].numpy().astype(np.int64)
        for key in ("q", "k_new", "v_new", "k_pool", "v_pool"):
            out[f"{n}.{key}"] = bits(c[key])
        # pool rows the store wrote (the rest of the pool is unchanged by construction)
        out[f"{n}.k_rows_after"] = bits(pool.k_buffer[0][c["out_cache_loc"]])
        out[f"{n}.v_rows_after"] = bits(pool.v_buffer[0][c["out_cache_loc"]])
        out[f"{n}.out_ref"] = bits(o)
        out[f"{n}.out_ref_fp32"] = o32.detach().numpy().astype(np.float32)
        assert not is_fp8 or pool.k_buffer[0].dtype == torch.uint8
    out["names"] = np.array(names)
    np.savez_compressed(GOLDEN / "attention.npz", **out)
    total = sum(v.nbytes for v in out.values())
    print(f"attention.npz: {len(names)} cases, {total / 1e6:.2f} MB raw")


def main():
    GOLDEN.mkdir(parents=True, exist_ok=True)
    torch.set_num_threads(8)
    gen_cast_tables()
    gen_kv_store()
    gen_attention()


if __name__ == "__main__":
    sys.exit(main())
