"""Tensor-level entry points over the C ABI (include/b200attn.h).

PyTorch is plumbing here: device memory, the current CUDA stream and dtype tags.
Every function launches hand-written sm_100a kernels from libb200attn.so on the
current stream and allocates nothing, so the calls are CUDA-graph capturable.
There is no CPU path: CPU tensors raise.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import torch

from . import _C

_DTYPE_CODE = {
    torch.bfloat16: _C.B200_BF16,
    torch.float16: _C.B200_F16,
    torch.float8_e4m3fn: _C.B200_FP8_E4M3,
    torch.float8_e5m2: _C.B200_FP8_E5M2,
}

KERNEL_AUTO, KERNEL_GENERIC, KERNEL_FAST = _C.KERNEL_AUTO, _C.KERNEL_GENERIC, _C.KERNEL_FAST

# launch counter: bench.py reports it as `gpu_launches`
launch_count = 0


def dtype_code(dtype: torch.dtype) -> int:
    try:
        return _DTYPE_CODE[dtype]
    except KeyError:
        raise ValueError(f"sglang_b200: unsupported dtype {dtype}") from None


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "sglang_b200 ops run only on CUDA tensors (no CPU fallback); got a tensor on " + str(t.device)
            )


def _index_kind(t: torch.Tensor, name: str) -> int:
    if t.dtype == torch.int64:
        return 1
    if t.dtype == torch.int32:
        return 0
    raise ValueError(f"{name} must be int32 or int64, got {t.dtype}")


def _f32_vec(t: Optional[torch.Tensor], n: int, name: str):
    if t is None:
        return None
    _need_cuda(t)
    if t.dtype != torch.float32 or t.shape != (n,) or not t.is_contiguous():
        raise ValueError(f"{name} must be a contiguous float32 tensor of shape ({n},)")
    return t.data_ptr()


def _f32_out(t: Optional[torch.Tensor], shape, name: str):
    if t is None:
        return None
    _need_cuda(t)
    if t.dtype != torch.float32 or tuple(t.shape) != tuple(shape) or not t.is_contiguous():
        raise ValueError(f"{name} must be a contiguous float32 tensor of shape {tuple(shape)}")
    return t.data_ptr()


_device_info_cache = {}


def device_info(device: Optional[torch.device] = None):
    """(sm_count, cc_major, cc_minor) of the current device; raises unless it is sm_100."""
    dev = torch.cuda.current_device() if device is None else torch.device(device).index or 0
    if dev not in _device_info_cache:
        sm, mj, mn = C.c_int(), C.c_int(), C.c_int()
        with torch.cuda.device(dev):
            _C.check(_C.lib.b200attn_device_info(C.byref(sm), C.byref(mj), C.byref(mn)), "device_info")
        _device_info_cache[dev] = (sm.value, mj.value, mn.value)
    return _device_info_cache[dev]


# --------------------------------------------------------------------------- KV store
def store_kv(
    k: torch.Tensor,
    v: torch.Tensor,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    loc: torch.Tensor,
    *,
    cache_dtype: Optional[torch.dtype] = None,
    k_scale: Optional[float] = None,
    v_scale: Optional[float] = None,
    size_limit: int = 0,
    skip_index: int = 0,
) -> None:
    """k_cache[loc[i]] = cast(k[i]), v_cache[loc[i]] = cast(v[i]).

    Mirrors `store_cache` (reference kernels/ops/kvcache/kvcache.py:57-122) plus the
    cast/scale rules of `MHATokenToKVPool.set_kv_buffer` (memory_pool.py:2339-2349):
    `cache_dtype` names the logical pool dtype when the buffers are uint8 views of FP8.
    """
    global launch_count
    _need_cuda(k, v, k_cache, v_cache, loc)
    T = k.shape[0]
    if T == 0:
        return
    k2 = k.reshape(T, -1)
    v2 = v.reshape(T, -1)
    if k2.stride(-1) != 1 or v2.stride(-1) != 1:
        k2, v2 = k2.contiguous(), v2.contiguous()
    kc = k_cache.view(k_cache.shape[0], -1)
    vc = v_cache.view(v_cache.shape[0], -1)
    if cache_dtype is None:
        cache_dtype = k_cache.dtype
    if loc.dim() != 1 or loc.shape[0] != T or loc.stride(0) != 1:
        raise ValueError("loc must be a contiguous 1-D tensor with one slot id per row of k")
    a = _C.StoreKVArgs(
        k=k2.data_ptr(), v=v2.data_ptr(), k_src_stride=k2.stride(0), v_src_stride=v2.stride(0),
        k_cache=kc.data_ptr(), v_cache=vc.data_ptr(), k_cache_stride=kc.stride(0), v_cache_stride=vc.stride(0),
        loc=loc.data_ptr(), loc_is_i64=_index_kind(loc, "loc"), num_tokens=T,
        k_row_elems=k2.shape[1], v_row_elems=v2.shape[1],
        src_dtype=dtype_code(k.dtype), dst_dtype=dtype_code(cache_dtype),
        k_scale=1.0 if k_scale is None else float(k_scale),
        v_scale=1.0 if v_scale is None else float(v_scale),
        size_limit=size_limit if size_limit > 0 else kc.shape[0], skip_index=skip_index,
    )
    if kc.shape[1] != k2.shape[1] or vc.shape[1] != v2.shape[1]:
        raise ValueError(f"row width mismatch: k {tuple(k2.shape)} vs cache {tuple(kc.shape)}")
    _C.check(_C.lib.b200attn_store_kv(C.byref(a), _stream()), "store_kv")
    launch_count += 1


def rope_store_kv(
    positions: torch.Tensor,       # [T] int32 / int64
    q: torch.Tensor,               # [T, Hq * D] or [T, Hq, D], rotated in place
    k: torch.Tensor,               # [T, Hkv * D] or [T, Hkv, D], rotated in place when write_k
    v: Optional[torch.Tensor],     # [T, Hkv * Dv] or [T, Hkv, Dv]
    cos_sin_cache: torch.Tensor,   # [max_position, rotary_dim] float32 or the model dtype
    head_dim: int,
    *,
    is_neox: bool = True,
    k_cache: Optional[torch.Tensor] = None,   # [slots, Hkv, D]; None = rotate only
    v_cache: Optional[torch.Tensor] = None,
    loc: Optional[torch.Tensor] = None,
    cache_dtype: Optional[torch.dtype] = None,
    k_scale: Optional[float] = None,
    v_scale: Optional[float] = None,
    native_rounding: bool = False,
    write_k: bool = True,
    skip_index: int = 0,
) -> None:
    """Fused `rotary_emb(positions, q, k)` + `set_kv_buffer(layer, loc, k, v)` in one launch.

    Mirrors RotaryEmbedding.forward_cuda with `fused_set_kv_buffer_arg` (reference
    srt/layers/rotary_embedding/base.py:363-385) followed by the cast/scale rules of
    `MHATokenToKVPool.set_kv_buffer`.  `native_rounding=True` reproduces `forward_native`
    (every product and sum rounded to the model dtype) bit for bit; the default computes each output
    in fp32 with one rounding, like the reference's CUDA kernel."""
    global launch_count
    T = positions.shape[0]
    if T == 0:
        return
    _need_cuda(positions, q, k, cos_sin_cache)
    q3, k3 = q.view(T, -1, head_dim), k.view(T, -1, head_dim)
    if q3.stride(2) != 1 or k3.stride(2) != 1 or q3.stride(1) != head_dim or k3.stride(1) != head_dim:
        raise ValueError("q / k heads must be contiguous ([T, H, D] with strides (*, D, 1))")
    if cos_sin_cache.dim() != 2 or not cos_sin_cache.is_contiguous():
        raise ValueError("cos_sin_cache must be a contiguous [max_position, rotary_dim] tensor")
    if cos_sin_cache.dtype == torch.float32:
        cdt = _C.CACHE_F32
    elif cos_sin_cache.dtype == q.dtype:
        cdt = _C.CACHE_MODEL_DTYPE
    else:
        raise ValueError(f"cos_sin_cache must be float32 or {q.dtype}, got {cos_sin_cache.dtype}")
    Hq, Hkv = q3.shape[1], k3.shape[1]
    a = _C.RopeStoreArgs(
        q=q3.data_ptr(), k=k3.data_ptr(), q_stride_t=q3.stride(0), k_stride_t=k3.stride(0),
        positions=positions.data_ptr(), positions_is_i64=_index_kind(positions, "positions"),
        cos_sin_cache=cos_sin_cache.data_ptr(), cache_dtype=cdt, max_position=cos_sin_cache.shape[0],
        rotary_dim=cos_sin_cache.shape[1], head_dim=head_dim, v_head_dim=head_dim,
        num_q_heads=Hq, num_kv_heads=Hkv, is_neox=1 if is_neox else 0,
        rounding=_C.ROPE_NATIVE if native_rounding else _C.ROPE_FP32, write_k=1 if write_k else 0,
        num_tokens=T, src_dtype=dtype_code(q.dtype), dst_dtype=dtype_code(q.dtype),
        k_scale=1.0 if k_scale is None else float(k_scale), v_scale=1.0 if v_scale is None else float(v_scale),
        size_limit=0, skip_index=skip_index,
    )
    if k_cache is not None:
        if v is None or v_cache is None or loc is None:
            raise ValueError("k_cache needs v, v_cache and loc as well")
        _need_cuda(v, k_cache, v_cache, loc)
        kc, vc = k_cache.view(k_cache.shape[0], -1), v_cache.view(v_cache.shape[0], -1)
        v3 = v.view(T, Hkv, -1)
        if v3.stride(2) != 1 or v3.stride(1) != v3.shape[2]:
            raise ValueError("v heads must be contiguous")
        if kc.shape[1] != Hkv * head_dim or vc.shape[1] != Hkv * v3.shape[2]:
            raise ValueError(f"row width mismatch: k/v {Hkv}x{head_dim}/{v3.shape[2]} vs cache {tuple(kc.shape)}")
        if loc.dim() != 1 or loc.shape[0] != T or loc.stride(0) != 1:
            raise ValueError("loc must be a contiguous 1-D tensor with one slot id per token")
        a.v, a.v_stride_t, a.v_head_dim = v3.data_ptr(), v3.stride(0), v3.shape[2]
        a.k_cache, a.v_cache = kc.data_ptr(), vc.data_ptr()
        a.k_cache_stride, a.v_cache_stride = kc.stride(0), vc.stride(0)
        a.loc, a.loc_is_i64 = loc.data_ptr(), _index_kind(loc, "loc")
        a.dst_dtype = dtype_code(cache_dtype or k_cache.dtype)
        a.size_limit = kc.shape[0]
    _C.check(_C.lib.b200attn_rope_store_kv(C.byref(a), _stream()), "rope_store_kv")
    launch_count += 1


# --------------------------------------------------------------------------- decode
@dataclass
class DecodeWorkspace:
    """Device buffers whose addresses are baked into CUDA graphs (allocated once)."""

    max_bs: int
    max_chunks: int
    num_ctas: int
    plan: torch.Tensor   # int32
    ws_o: torch.Tensor   # float32 [max_chunks * Hq * Dv]
    ws_lse: torch.Tensor  # float32 [max_chunks * Hq]

    @staticmethod
    def create(max_bs: int, num_q_heads: int, head_dim_v: int, device, *, num_ctas: int,
               waves: int = 16) -> "DecodeWorkspace":
        max_chunks = num_ctas * waves * 2 + max_bs + 64
        n_plan = _C.lib.b200attn_decode_plan_ints(max_bs, max_chunks)
        return DecodeWorkspace(
            max_bs=max_bs, max_chunks=max_chunks, num_ctas=num_ctas,
            plan=torch.zeros(n_plan, dtype=torch.int32, device=device),
            ws_o=torch.empty(max_chunks * num_q_heads * head_dim_v, dtype=torch.float32, device=device),
            ws_lse=torch.empty(max_chunks * num_q_heads, dtype=torch.float32, device=device),
        )


def decode_plan(ws: DecodeWorkspace, seq_lens: torch.Tensor, num_kv_heads: int, *, waves: int = 16,
                min_chunk_tokens: int = 512, fixed_chunk_tokens: int = 0,
                items_per_chunk: Optional[int] = None, window_keys: int = 0) -> None:
    """Per-forward split-KV work list, built on the device (no host sync).

    `items_per_chunk` (see `decode_items_per_chunk`) replaces `num_kv_heads` as the grid hint when the
    head-group kernel will run: an item then covers 4 kv heads, so the smallest useful chunk shrinks by
    the same factor.  `window_keys` > 0 plans only the last `window_keys` tokens of every request (sliding
    window; pass the same value to `decode_attention`)."""
    global launch_count
    _need_cuda(seq_lens, ws.plan)
    if items_per_chunk is not None and 0 < items_per_chunk < num_kv_heads:
        min_chunk_tokens = max(64, (min_chunk_tokens * items_per_chunk // num_kv_heads) // 16 * 16)
        num_kv_heads = items_per_chunk
    a = _C.DecodePlanArgs(
        seq_lens=seq_lens.data_ptr(), seq_lens_is_i64=_index_kind(seq_lens, "seq_lens"),
        batch_size=seq_lens.shape[0], max_bs=ws.max_bs, max_chunks=ws.max_chunks,
        num_kv_heads=num_kv_heads, num_ctas=ws.num_ctas, waves=waves,
        min_chunk_tokens=min_chunk_tokens, fixed_chunk_tokens=fixed_chunk_tokens, plan=ws.plan.data_ptr(),
        window_keys=max(0, int(window_keys)),
    )
    _C.check(_C.lib.b200attn_decode_plan(C.byref(a), _stream()), "decode_plan")
    launch_count += 1


def decode_attention(
    q: torch.Tensor,            # [bs, Hq, Dk]
    o: torch.Tensor,            # [bs, Hq, Dv]
    k_cache: torch.Tensor,      # [slots, Hkv, Dk]
    v_cache: torch.Tensor,      # [slots, Hkv, Dv]
    req_to_token: torch.Tensor,
    req_pool_indices: torch.Tensor,
    seq_lens: torch.Tensor,
    ws: DecodeWorkspace,
    sm_scale: float,
    *,
    kv_dtype: Optional[torch.dtype] = None,
    k_scale: float = 1.0,
    v_scale: float = 1.0,
    logit_cap: float = 0.0,
    kernel: int = KERNEL_AUTO,
    attn_events=None,
    window_keys: int = 0,
    sinks: Optional[torch.Tensor] = None,
    lse: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """`window_keys` > 0: sliding window over the last `window_keys` tokens (the plan must have been built with
    the same value); `sinks`: float32 [Hq] attention-sink logits; `lse`: float32 [bs, Hq] natural-log LSE output.
    `attn_events`: optional (start, stop) torch.cuda.Event pair (already recorded once, so the
    handles exist) that the library re-records around the attention kernel alone -- bench.py's
    live per-kernel timing."""
    global launch_count
    _need_cuda(q, o, k_cache, v_cache, req_to_token, req_pool_indices, seq_lens)
    if q.dim() != 3 or o.dim() != 3 or k_cache.dim() != 3 or v_cache.dim() != 3:
        raise ValueError("decode_attention expects q/o [bs,H,D] and caches [slots,Hkv,D]")
    if q.stride(2) != 1 or o.stride(2) != 1 or k_cache.stride(2) != 1 or v_cache.stride(2) != 1:
        raise ValueError("innermost (head_dim) stride must be 1")
    if req_to_token.dtype != torch.int32:
        raise ValueError("req_to_token must be int32")
    bs, Hq, Dk = q.shape
    Hkv, Dv = k_cache.shape[1], v_cache.shape[2]
    a = _C.DecodeArgs(
        q=q.data_ptr(), o=o.data_ptr(),
        q_stride_t=q.stride(0), q_stride_h=q.stride(1), o_stride_t=o.stride(0), o_stride_h=o.stride(1),
        k_cache=k_cache.data_ptr(), v_cache=v_cache.data_ptr(),
        k_stride_s=k_cache.stride(0), k_stride_h=k_cache.stride(1),
        v_stride_s=v_cache.stride(0), v_stride_h=v_cache.stride(1),
        req_to_token=req_to_token.data_ptr(), req_to_token_stride=req_to_token.stride(0),
        req_pool_indices=req_pool_indices.data_ptr(),
        req_pool_indices_is_i64=_index_kind(req_pool_indices, "req_pool_indices"),
        seq_lens=seq_lens.data_ptr(), seq_lens_is_i64=_index_kind(seq_lens, "seq_lens"),
        batch_size=bs, num_q_heads=Hq, num_kv_heads=Hkv, head_dim_qk=Dk, head_dim_v=Dv,
        q_dtype=dtype_code(q.dtype), kv_dtype=dtype_code(kv_dtype or k_cache.dtype),
        sm_scale=float(sm_scale), k_scale=float(k_scale), v_scale=float(v_scale), logit_cap=float(logit_cap),
        plan=ws.plan.data_ptr(), max_bs=ws.max_bs, max_chunks=ws.max_chunks,
        ws_o=ws.ws_o.data_ptr(), ws_lse=ws.ws_lse.data_ptr(), num_ctas=ws.num_ctas, kernel=kernel,
        ev_attn_start=attn_events[0].cuda_event if attn_events else None,
        ev_attn_stop=attn_events[1].cuda_event if attn_events else None,
        window_keys=max(0, int(window_keys)), sinks=_f32_vec(sinks, Hq, "sinks"),
        lse=_f32_out(lse, (bs, Hq), "lse"),
    )
    if ws.ws_o.numel() < ws.max_chunks * Hq * Dv or ws.ws_lse.numel() < ws.max_chunks * Hq:
        raise ValueError("decode workspace too small for this head configuration")
    _C.check(_C.lib.b200attn_decode(C.byref(a), _stream()), "decode")
    launch_count += 2  # attention + split merge
    return o


def build_draft_table(table: torch.Tensor, req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                      seq_lens: torch.Tensor, topk: int, num_steps: int, page_size: int = 1) -> torch.Tensor:
    """EAGLE tree drafting: fill `table` [num_seqs * topk, W] (int32) with one page-table row per (request, branch):
    the request's prefix followed by that branch's drafted slots (reference
    kernels/ops/speculative/cache_locs.py:56-130), so that step i of the draft is a plain decode over `table` with
    lengths `seq_lens + i + 1`.  Graph-recordable (lengths are read on the device)."""
    global launch_count
    _need_cuda(table, req_to_token, req_pool_indices, seq_lens)
    n = seq_lens.shape[0]
    if table.dtype != torch.int32 or table.dim() != 2 or table.stride(1) != 1 or table.shape[0] < n * topk:
        raise ValueError("table must be int32 [>= num_seqs * topk, W] with contiguous rows")
    if req_to_token.dtype != torch.int32 or req_to_token.stride(1) != 1:
        raise ValueError("req_to_token must be int32 with contiguous rows")
    a = _C.DraftTableArgs(
        req_to_token=req_to_token.data_ptr(), req_to_token_stride=req_to_token.stride(0),
        req_pool_indices=req_pool_indices.data_ptr(),
        req_pool_indices_is_i64=_index_kind(req_pool_indices, "req_pool_indices"),
        seq_lens=seq_lens.data_ptr(), seq_lens_is_i64=_index_kind(seq_lens, "seq_lens"),
        num_seqs=n, topk=int(topk), num_steps=int(num_steps), page_size=max(1, int(page_size)),
        table=table.data_ptr(), table_stride=table.stride(0),
    )
    _C.check(_C.lib.b200attn_draft_table(C.byref(a), _stream()), "draft_table")
    launch_count += 1
    return table


def decode_fast_supported(head_dim_qk: int, head_dim_v: int, group: int, q_dtype, kv_dtype) -> bool:
    return bool(_C.lib.b200attn_decode_fast_supported(head_dim_qk, head_dim_v, group, dtype_code(q_dtype),
                                                      dtype_code(kv_dtype)))


def decode_items_per_chunk(num_q_heads: int, num_kv_heads: int, head_dim_qk: int, head_dim_v: int,
                           q_dtype, kv_dtype) -> int:
    """Work items per split chunk (kv heads, or kv-head groups of 4 for the head-group kernel): the
    `num_kv_heads` grid hint `decode_plan` wants."""
    return int(_C.lib.b200attn_decode_items_per_chunk(num_q_heads, num_kv_heads, head_dim_qk, head_dim_v,
                                                      dtype_code(q_dtype), dtype_code(kv_dtype)))


def decode_ctas_per_sm(head_dim: int, kv_dtype) -> int:
    return int(_C.lib.b200attn_decode_ctas_per_sm(head_dim, dtype_code(kv_dtype)))


# --------------------------------------------------------------------------- extend
@dataclass
class ExtendWorkspace:
    max_bs: int
    max_tiles: int
    num_ctas: int
    plan: torch.Tensor
    block_q: int = 0          # what the current plan was built with (set by extend_plan)

    @staticmethod
    def create(max_bs: int, max_tiles: int, device, *, num_ctas: int) -> "ExtendWorkspace":
        n_plan = _C.lib.b200attn_extend_plan_ints(max_bs, max_tiles)
        return ExtendWorkspace(max_bs=max_bs, max_tiles=max_tiles, num_ctas=num_ctas,
                               plan=torch.zeros(n_plan, dtype=torch.int32, device=device))


def extend_block_q(head_dim_qk: int, head_dim_v: int, group: int, q_dtype, kv_dtype,
                   kernel: int = KERNEL_AUTO) -> int:
    """Query tokens per work item of the kernel that will run; pass it to `extend_plan`."""
    return int(_C.lib.b200attn_extend_block_q(head_dim_qk, head_dim_v, group, dtype_code(q_dtype),
                                              dtype_code(kv_dtype), kernel))


def extend_block_q_one_tile(head_dim_qk: int, head_dim_v: int, group: int, q_dtype, kv_dtype,
                             kernel: int = KERNEL_AUTO) -> int:
    """block_q of the tcgen05 kernel's one-tile instance (128 / group tokens per work item; pass it to `extend_plan`
    when no request extends by more than that many tokens), or 0 when the generic kernel will run."""
    if kernel == KERNEL_GENERIC or not extend_fast_supported(head_dim_qk, head_dim_v, group, q_dtype, kv_dtype):
        return 0
    return 128 // group


def extend_fast_supported(head_dim_qk: int, head_dim_v: int, group: int, q_dtype, kv_dtype) -> bool:
    return bool(_C.lib.b200attn_extend_fast_supported(head_dim_qk, head_dim_v, group, dtype_code(q_dtype),
                                                      dtype_code(kv_dtype)))


def extend_plan(ws: ExtendWorkspace, extend_seq_lens: torch.Tensor, block_q: int, *,
                page_size: int = 1, req_to_token: Optional[torch.Tensor] = None,
                req_pool_indices: Optional[torch.Tensor] = None,
                extend_prefix_lens: Optional[torch.Tensor] = None) -> None:
    """Tile list for `extend_attention`.  With `page_size` >= 16 and the page table given, the plan also verifies on
    the device that every page is a run of consecutive pool rows; `extend_attention(page_size=...)` then moves K/V by
    TMA page boxes, otherwise (or when the check fails) row by row -- same result either way."""
    global launch_count
    _need_cuda(extend_seq_lens, ws.plan)
    a = _C.ExtendPlanArgs(
        extend_seq_lens=extend_seq_lens.data_ptr(), lens_is_i64=_index_kind(extend_seq_lens, "extend_seq_lens"),
        batch_size=extend_seq_lens.shape[0], max_bs=ws.max_bs, max_tiles=ws.max_tiles, block_q=block_q,
        plan=ws.plan.data_ptr(),
    )
    check = page_size >= 16 and req_to_token is not None and extend_seq_lens.shape[0] > 0
    if check:
        if req_pool_indices is None or extend_prefix_lens is None:
            raise ValueError("extend_plan: the page check needs req_pool_indices and extend_prefix_lens")
        _need_cuda(req_to_token, req_pool_indices, extend_prefix_lens)
        if req_to_token.dtype != torch.int32 or req_to_token.stride(1) != 1:
            raise ValueError("req_to_token must be int32 with contiguous rows")
        if extend_prefix_lens.dtype != extend_seq_lens.dtype:
            raise ValueError("extend_prefix_lens and extend_seq_lens must share a dtype")
        a.req_to_token = req_to_token.data_ptr()
        a.req_to_token_stride = req_to_token.stride(0)
        a.req_pool_indices = req_pool_indices.data_ptr()
        a.req_pool_indices_is_i64 = _index_kind(req_pool_indices, "req_pool_indices")
        a.extend_prefix_lens = extend_prefix_lens.data_ptr()
        a.page_size = int(page_size)
    _C.check(_C.lib.b200attn_extend_plan(C.byref(a), _stream()), "extend_plan")
    ws.block_q = int(block_q)
    launch_count += 2 if check and (page_size & (page_size - 1)) == 0 else 1


def extend_attention(
    q: torch.Tensor,            # [T, Hq, Dk]
    o: torch.Tensor,            # [T, Hq, Dv]
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    req_to_token: torch.Tensor,
    req_pool_indices: torch.Tensor,
    extend_prefix_lens: torch.Tensor,
    extend_seq_lens: torch.Tensor,
    ws: ExtendWorkspace,
    sm_scale: float,
    *,
    kv_dtype: Optional[torch.dtype] = None,
    k_scale: float = 1.0,
    v_scale: float = 1.0,
    logit_cap: float = 0.0,
    causal: bool = True,
    kernel: int = KERNEL_AUTO,
    lse: Optional[torch.Tensor] = None,
    window_keys: int = 0,
    sinks: Optional[torch.Tensor] = None,
    custom_mask: Optional[torch.Tensor] = None,
    mask_indptr: Optional[torch.Tensor] = None,
    mask_skip_prefix: bool = True,
    page_size: int = 1,
    num_splits: int = 1,
    split_ws: Optional[torch.Tensor] = None,
    num_tiles_hint: int = 0,
) -> torch.Tensor:
    """`num_splits` > 1 (one-tile plans only, see `extend_pick_splits`): split-KV -- every work item's kv range is cut
    into up to `num_splits` parts that run as separate CTAs, `split_ws` (float32, >= `extend_split_ws_floats(...)`)
    holds their partials and a merge launch follows.  `num_tiles_hint`: the exact tile count when the host knows it
    (one per request for speculative batches) -- batches of more items than SMs are then split in the tail only.
    `page_size`: KV page size of the pool (pages are page_size-aligned runs of consecutive slots).
    `window_keys` > 0: query at position x attends kv positions (x - window_keys, x].  `sinks`: float32 [Hq].
    `custom_mask` (uint8 / bool, flat) + `mask_indptr` (int64 [bs + 1]): request r, query i, kv position x is
    visible iff custom_mask[mask_indptr[r] + i * (P_r + E_r) + x]; it replaces the causal rule on the new tokens
    and, unless `mask_skip_prefix`, also masks the prefix (reference extend_attention.py:378-390,525-539)."""
    global launch_count
    _need_cuda(q, o, k_cache, v_cache, req_to_token, req_pool_indices, extend_prefix_lens, extend_seq_lens)
    if q.dim() != 3 or o.dim() != 3 or k_cache.dim() != 3 or v_cache.dim() != 3:
        raise ValueError("extend_attention expects q/o [T,H,D] and caches [slots,Hkv,D]")
    if q.stride(2) != 1 or o.stride(2) != 1 or k_cache.stride(2) != 1 or v_cache.stride(2) != 1:
        raise ValueError("innermost (head_dim) stride must be 1")
    if req_to_token.dtype != torch.int32:
        raise ValueError("req_to_token must be int32")
    if extend_prefix_lens.dtype != extend_seq_lens.dtype:
        raise ValueError("extend_prefix_lens and extend_seq_lens must share a dtype")
    T, Hq, Dk = q.shape
    Hkv, Dv = k_cache.shape[1], v_cache.shape[2]
    a = _C.ExtendArgs(
        q=q.data_ptr(), o=o.data_ptr(),
        q_stride_t=q.stride(0), q_stride_h=q.stride(1), o_stride_t=o.stride(0), o_stride_h=o.stride(1),
        k_cache=k_cache.data_ptr(), v_cache=v_cache.data_ptr(),
        k_stride_s=k_cache.stride(0), k_stride_h=k_cache.stride(1),
        v_stride_s=v_cache.stride(0), v_stride_h=v_cache.stride(1),
        req_to_token=req_to_token.data_ptr(), req_to_token_stride=req_to_token.stride(0),
        req_pool_indices=req_pool_indices.data_ptr(),
        req_pool_indices_is_i64=_index_kind(req_pool_indices, "req_pool_indices"),
        extend_prefix_lens=extend_prefix_lens.data_ptr(), extend_seq_lens=extend_seq_lens.data_ptr(),
        lens_is_i64=_index_kind(extend_seq_lens, "extend_seq_lens"),
        batch_size=extend_seq_lens.shape[0], num_q_tokens=T,
        num_q_heads=Hq, num_kv_heads=Hkv, head_dim_qk=Dk, head_dim_v=Dv,
        q_dtype=dtype_code(q.dtype), kv_dtype=dtype_code(kv_dtype or k_cache.dtype),
        sm_scale=float(sm_scale), k_scale=float(k_scale), v_scale=float(v_scale), logit_cap=float(logit_cap),
        causal=1 if causal else 0, plan=ws.plan.data_ptr(), max_bs=ws.max_bs, max_tiles=ws.max_tiles,
        num_ctas=ws.num_ctas, kernel=kernel, lse=_f32_out(lse, (T, Hq), "lse"),
        pool_slots=k_cache.shape[0],
        window_keys=max(0, int(window_keys)), sinks=_f32_vec(sinks, Hq, "sinks"),
        mask_skip_prefix=1 if mask_skip_prefix else 0,
        block_q=ws.block_q, page_size=max(1, int(page_size)),
    )
    if custom_mask is not None:
        if mask_indptr is None:
            raise ValueError("custom_mask needs mask_indptr")
        _need_cuda(custom_mask, mask_indptr)
        if custom_mask.dtype not in (torch.uint8, torch.bool) or not custom_mask.is_contiguous():
            raise ValueError("custom_mask must be a contiguous uint8 / bool tensor")
        if mask_indptr.dtype != torch.int64 or mask_indptr.shape != (extend_seq_lens.shape[0] + 1,):
            raise ValueError("mask_indptr must be int64 [bs + 1]")
        a.custom_mask, a.mask_indptr = custom_mask.data_ptr(), mask_indptr.data_ptr()
    if num_splits > 1:
        if split_ws is None:
            raise ValueError("num_splits > 1 needs split_ws")
        _need_cuda(split_ws)
        if split_ws.dtype != torch.float32 or not split_ws.is_contiguous():
            raise ValueError("split_ws must be a contiguous float32 tensor")
        a.num_splits, a.split_ws, a.split_ws_floats = int(num_splits), split_ws.data_ptr(), split_ws.numel()
        a.num_tiles_hint = max(0, int(num_tiles_hint))
    _C.check(_C.lib.b200attn_extend(C.byref(a), _stream()), "extend")
    launch_count += 2 if num_splits > 1 else 1
    return o


def extend_pick_splits(items: int, num_sms: int, kv_tokens_hint: int = 0) -> int:
    """Split count for `items` one-tile work items (requests x kv heads) of about `kv_tokens_hint` kv tokens each
    (0 = unknown) on `num_sms` SMs; 1 = no split."""
    return int(_C.lib.b200attn_extend_pick_splits(int(items), int(num_sms), max(0, int(kv_tokens_hint))))


def extend_split_ws_floats(num_splits: int, num_q_tokens: int, num_q_heads: int, head_dim_v: int) -> int:
    return int(_C.lib.b200attn_extend_split_ws_floats(int(num_splits), int(num_q_tokens), int(num_q_heads),
                                                      int(head_dim_v)))


# --------------------------------------------------------------------------- merge of partial states
def merge_state(o_a: torch.Tensor, lse_a: torch.Tensor, o_b: torch.Tensor, lse_b: torch.Tensor,
                out: Optional[torch.Tensor] = None, out_lse: Optional[torch.Tensor] = None):
    """Merge two partial attention states [T, H, D] + natural-log LSE [T, H] (chunked prefix / cascade).

    Mirrors `merge_state_v2` / `merge_state_triton` (reference kernels/ops/attention/merge_state.py:69-100):
    returns (out, out_lse); an LSE of +inf counts as an empty state."""
    global launch_count
    _need_cuda(o_a, lse_a, o_b, lse_b, out, out_lse)
    if o_a.dim() != 3 or o_a.shape != o_b.shape or o_a.dtype != o_b.dtype:
        raise ValueError("merge_state expects two [T, H, D] tensors of the same shape and dtype")
    T, H, D = o_a.shape
    if out is None:
        out = torch.empty_like(o_a)
    if out_lse is None:
        out_lse = torch.empty((T, H), dtype=torch.float32, device=o_a.device)
    for t, name in ((lse_a, "lse_a"), (lse_b, "lse_b"), (out_lse, "out_lse")):
        if t.dtype != torch.float32 or tuple(t.shape) != (T, H) or not t.is_contiguous():
            raise ValueError(f"{name} must be a contiguous float32 [T, H] tensor")
    for t in (o_a, o_b, out):
        if t.stride(2) != 1:
            raise ValueError("innermost (head_dim) stride must be 1")
    if T == 0:
        return out, out_lse
    a = _C.MergeStateArgs(
        o_a=o_a.data_ptr(), lse_a=lse_a.data_ptr(), o_b=o_b.data_ptr(), lse_b=lse_b.data_ptr(),
        o=out.data_ptr(), lse=out_lse.data_ptr(), num_tokens=T, num_heads=H, head_dim=D,
        dtype=dtype_code(o_a.dtype),
        oa_stride_t=o_a.stride(0), oa_stride_h=o_a.stride(1), ob_stride_t=o_b.stride(0), ob_stride_h=o_b.stride(1),
        o_stride_t=out.stride(0), o_stride_h=out.stride(1),
    )
    _C.check(_C.lib.b200attn_merge_state(C.byref(a), _stream()), "merge_state")
    launch_count += 1
    return out, out_lse
