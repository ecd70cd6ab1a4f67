"""ORACLE (test infrastructure): FP32-accumulated restatement of the reference CPU path.

Follows, function by function (paths relative to the reference's python/sglang/srt/):
  decode_ref   layers/attention/torch_native_backend.py:176-277  (_run_sdpa_forward_decode)
               + :338-402 (forward_decode: store first, then attend through the page table)
  extend_ref   layers/attention/torch_native_backend.py:61-174   (_run_sdpa_forward_extend)
               + :279-336 (forward_extend); the reference pads Q to seq_len rows, runs causal
               SDPA over seq_len x seq_len and keeps rows prefix: (:128-134, :172) -- here the
               same rows are computed directly with the equivalent mask  kv_pos <= prefix + i
  store_kv_ref mem_cache/memory_pool.py:2305-2381 (set_kv_buffer) -> :141-192 (scatter);
               byte work delegated to the C restatement oracle/kv_store_ref.c
  FP32 math    test kit reference python/sglang/test/kits/attention_unittest/
               attention_methods/dense_attention.py:795-848 (softmax(q k^T * scaling) v in float32)

Pinned by tests/test_oracle.py against tests/golden/*.npz, which oracle/gen_golden.py produced
by running the reference itself (TorchNativeAttnBackend + MHATokenToKVPool) on CPU.  The options the
torch_native backend does not have (attention sinks, logit cap, custom masks, LSE output, merge_state) and
its sliding window are pinned by tests/golden/features.npz, which oracle/gen_golden_features.py produced
by running the reference's own Triton kernels (extend_attention_fwd, decode_attention_fwd,
merge_state_triton) under TRITON_INTERPRET=1 on CPU, plus TorchNativeAttnBackend with sliding windows.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Optional

import numpy as np
import torch

_REF_CODE = {torch.bfloat16: 0, torch.float16: 1, torch.float8_e4m3fn: 2, torch.float8_e5m2: 3}
_LIB: Optional[C.CDLL] = None


def _lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        path = Path(__file__).resolve().parent / "_build" / "liboracle_kv.so"
        if not path.exists():
            from oracle.build_oracle import build_oracle

            build_oracle()
        lib = C.CDLL(str(path))
        lib.ref_store_kv.restype = C.c_int64
        lib.ref_store_kv.argtypes = [
            C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
            C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_float, C.c_float,
            C.c_int64, C.c_int64,
        ]
        lib.ref_cast_table.restype = None
        lib.ref_cast_table.argtypes = [C.c_int, C.c_int, C.c_void_p]
        _LIB = lib
    return _LIB


def cast_table(src_dtype: torch.dtype, dst_dtype: torch.dtype) -> np.ndarray:
    """uint8[65536]: FP8 byte for every 16-bit source pattern (C restatement)."""
    out = np.zeros(65536, dtype=np.uint8)
    _lib().ref_cast_table(_REF_CODE[src_dtype], _REF_CODE[dst_dtype], out.ctypes.data)
    return out


def store_kv_ref(k: torch.Tensor, v: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                 loc: torch.Tensor, *, cache_dtype: torch.dtype, k_scale: Optional[float] = None,
                 v_scale: Optional[float] = None, skip_index: int = -1) -> None:
    """In-place scatter into CPU pool buffers (uint8 for FP8 pools, like `store_dtype`)."""
    assert not k.is_cuda and not k_cache.is_cuda
    T = k.shape[0]
    k2 = k.reshape(T, -1).contiguous()
    v2 = v.reshape(T, -1).contiguous()
    kc = k_cache.view(k_cache.shape[0], -1)
    vc = v_cache.view(v_cache.shape[0], -1)
    assert kc.is_contiguous() and vc.is_contiguous()
    loc64 = loc.to(torch.int64).contiguous()
    same = k.dtype == cache_dtype
    rc = _lib().ref_store_kv(
        k2.data_ptr(), v2.data_ptr(), k2.stride(0), v2.stride(0), kc.data_ptr(), vc.data_ptr(),
        kc.stride(0), vc.stride(0), loc64.data_ptr(), T, k2.shape[1], v2.shape[1],
        _REF_CODE[k.dtype], _REF_CODE[cache_dtype],
        1.0 if (same or k_scale is None) else float(k_scale),
        1.0 if (same or v_scale is None) else float(v_scale),
        kc.shape[0], skip_index,
    )
    if rc < 0:
        raise IndexError(f"store_kv_ref: loc[{-rc - 1}] out of range")


def _dequant(cache: torch.Tensor, kv_dtype: torch.dtype, scale: float) -> torch.Tensor:
    if cache.dtype == torch.uint8 and kv_dtype in (torch.float8_e4m3fn, torch.float8_e5m2):
        cache = cache.view(kv_dtype)
    out = cache.to(torch.float32)
    return out * scale if scale != 1.0 else out


def _attend(q: torch.Tensor, keys: torch.Tensor, values: torch.Tensor, scaling: float, group: int,
            mask: Optional[torch.Tensor], logit_cap: float, sinks: Optional[torch.Tensor] = None,
            want_lse: bool = False):
    """q [n_q, Hq, Dk], keys [n_kv, Hkv, Dk], values [n_kv, Hkv, Dv] (all float32) -> [n_q, Hq, Dv].

    `sinks` [Hq]: one extra logit per head that joins the softmax denominator but contributes no value
    (reference kernels/ops/attention/extend_attention.py:633-635).  `want_lse`: also return the natural-log
    LSE [n_q, Hq] (sink included, as the reference stores it at :637-642)."""
    k = keys.repeat_interleave(group, dim=1) if group > 1 else keys   # enable_gqa
    v = values.repeat_interleave(group, dim=1) if group > 1 else values
    s = torch.einsum("qhd,khd->hqk", q, k) * scaling
    if logit_cap > 0:
        s = logit_cap * torch.tanh(s / logit_cap)
    if mask is not None:
        s = s.masked_fill(~mask.unsqueeze(0), float("-inf"))
    if sinks is not None:
        sink_col = sinks.to(torch.float32).view(-1, 1, 1).expand(-1, s.shape[1], 1)
        s_all = torch.cat([s, sink_col], dim=-1)
        p = torch.softmax(s_all, dim=-1)[..., :-1]
        lse = torch.logsumexp(s_all, dim=-1)
    else:
        p = torch.softmax(s, dim=-1)
        lse = torch.logsumexp(s, dim=-1)
    out = torch.einsum("hqk,khd->qhd", p, v)
    return (out, lse.transpose(0, 1).contiguous()) if want_lse else out


def decode_ref(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, req_to_token: torch.Tensor,
               req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, scaling: float, *,
               kv_dtype: Optional[torch.dtype] = None, k_scale: float = 1.0, v_scale: float = 1.0,
               logit_cap: float = 0.0, window_keys: int = 0, sinks: Optional[torch.Tensor] = None,
               want_lse: bool = False):
    """q [bs, Hq, Dk] -> float32 [bs, Hq, Dv]; one query token per request.

    `window_keys` > 0: only the last `window_keys` tokens are attended -- torch_native's sliding-window mask
    k_pos >= q_pos - W (torch_native_backend.py:36-48, :251-258) with window_keys = W + 1."""
    kv_dtype = kv_dtype or k_cache.dtype
    Hq, Hkv = q.shape[1], k_cache.shape[1]
    group = Hq // Hkv
    qf = q.to(torch.float32)
    out = torch.empty(q.shape[0], Hq, v_cache.shape[2], dtype=torch.float32)
    lse = torch.empty(q.shape[0], Hq, dtype=torch.float32)
    for i in range(q.shape[0]):
        L = int(seq_lens[i])
        lo = max(0, L - window_keys) if window_keys > 0 else 0
        tok = req_to_token[int(req_pool_indices[i]), lo:L].long()
        keys = _dequant(k_cache[tok], kv_dtype, k_scale)
        vals = _dequant(v_cache[tok], kv_dtype, v_scale)
        o_i, l_i = _attend(qf[i : i + 1], keys, vals, scaling, group, None, logit_cap, sinks, True)
        out[i], lse[i] = o_i[0], l_i[0]
    return (out, lse) if want_lse else out


def extend_ref(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, req_to_token: torch.Tensor,
               req_pool_indices: torch.Tensor, extend_prefix_lens: torch.Tensor,
               extend_seq_lens: torch.Tensor, scaling: float, *, kv_dtype: Optional[torch.dtype] = None,
               k_scale: float = 1.0, v_scale: float = 1.0, logit_cap: float = 0.0,
               causal: bool = True, window_keys: int = 0, sinks: Optional[torch.Tensor] = None,
               custom_mask: Optional[torch.Tensor] = None, mask_indptr: Optional[torch.Tensor] = None,
               mask_skip_prefix: bool = True, want_lse: bool = False):
    """q [sum E_r, Hq, Dk] -> float32 [sum E_r, Hq, Dv]; the new tokens are already in the pool.

    `window_keys`: query at position x sees kv positions (x - window_keys, x] (torch_native_backend.py:150-157
    with sliding_window_size = window_keys - 1; Triton extend_attention.py:391-398, :550-555).
    `custom_mask` (flat uint8/bool) + `mask_indptr` [bs+1]: request r's block is an [E_r, P_r + E_r] matrix;
    it replaces the causal rule on the new tokens and, unless `mask_skip_prefix`, also masks the prefix
    (extend_attention.py:378-390, :525-539)."""
    kv_dtype = kv_dtype or k_cache.dtype
    Hq, Hkv = q.shape[1], k_cache.shape[1]
    group = Hq // Hkv
    qf = q.to(torch.float32)
    out = torch.zeros(q.shape[0], Hq, v_cache.shape[2], dtype=torch.float32)
    lse = torch.zeros(q.shape[0], Hq, dtype=torch.float32)
    start = 0
    for i in range(extend_seq_lens.shape[0]):
        P, E = int(extend_prefix_lens[i]), int(extend_seq_lens[i])
        if E == 0:
            continue
        tok = req_to_token[int(req_pool_indices[i]), : P + E].long()
        keys = _dequant(k_cache[tok], kv_dtype, k_scale)
        vals = _dequant(v_cache[tok], kv_dtype, v_scale)
        q_pos = torch.arange(P, P + E).unsqueeze(1)
        k_pos = torch.arange(P + E).unsqueeze(0)
        mask = None
        if custom_mask is not None:
            m0 = int(mask_indptr[i])
            mask = custom_mask[m0 : m0 + E * (P + E)].view(E, P + E).to(torch.bool).clone()
            if mask_skip_prefix:
                mask[:, :P] = True
        elif causal:
            mask = k_pos <= q_pos
        if window_keys > 0:
            win = k_pos > q_pos - window_keys
            mask = win if mask is None else (mask & win)
        o_i, l_i = _attend(qf[start : start + E], keys, vals, scaling, group, mask, logit_cap, sinks, True)
        out[start : start + E], lse[start : start + E] = o_i, l_i
        start += E
    return (out, lse) if want_lse else out


def merge_state_ref(o_a: torch.Tensor, lse_a: torch.Tensor, o_b: torch.Tensor, lse_b: torch.Tensor):
    """Float32 restatement of merge_state_kernel (reference kernels/ops/attention/merge_state.py:9-66,
    kernels/aot/csrc/attention/merge_attn_states.cu:60-120): +inf LSE counts as -inf."""
    la = torch.where(lse_a == float("inf"), torch.full_like(lse_a, float("-inf")), lse_a).float()
    lb = torch.where(lse_b == float("inf"), torch.full_like(lse_b, float("-inf")), lse_b).float()
    mx = torch.maximum(la, lb)
    ea, eb = torch.exp(la - mx), torch.exp(lb - mx)
    se = ea + eb
    out = o_a.float() * (ea / se).unsqueeze(-1) + o_b.float() * (eb / se).unsqueeze(-1)
    return out, torch.log(se) + mx
