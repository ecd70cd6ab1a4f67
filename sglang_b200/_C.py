"""ctypes binding of the C ABI in include/b200attn.h.

There is no CPU fallback: if the shared library is missing or does not load, importing
this module raises, and every op in sglang_b200.ops fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("B200ATTN_LIB", _PKG_DIR / "lib" / "libb200attn.so"))

B200_BF16, B200_F16, B200_FP8_E4M3, B200_FP8_E5M2 = 0, 1, 2, 3
KERNEL_AUTO, KERNEL_GENERIC, KERNEL_FAST = 0, 1, 2
ABI_VERSION = 8

i32, i64, f32, vp = C.c_int32, C.c_int64, C.c_float, C.c_void_p


class StoreKVArgs(C.Structure):
    _fields_ = [
        ("k", vp), ("v", vp), ("k_src_stride", i64), ("v_src_stride", i64),
        ("k_cache", vp), ("v_cache", vp), ("k_cache_stride", i64), ("v_cache_stride", i64),
        ("loc", vp), ("loc_is_i64", i32), ("num_tokens", i64),
        ("k_row_elems", i32), ("v_row_elems", i32), ("src_dtype", i32), ("dst_dtype", i32),
        ("k_scale", f32), ("v_scale", f32), ("size_limit", i64), ("skip_index", i64),
    ]


class RopeStoreArgs(C.Structure):
    _fields_ = [
        ("q", vp), ("k", vp), ("v", vp), ("q_stride_t", i64), ("k_stride_t", i64), ("v_stride_t", i64),
        ("positions", vp), ("positions_is_i64", i32), ("cos_sin_cache", vp), ("cache_dtype", i32),
        ("max_position", i64), ("rotary_dim", i32), ("head_dim", i32), ("v_head_dim", i32),
        ("num_q_heads", i32), ("num_kv_heads", i32), ("is_neox", i32), ("rounding", i32), ("write_k", i32),
        ("k_cache", vp), ("v_cache", vp), ("k_cache_stride", i64), ("v_cache_stride", i64),
        ("loc", vp), ("loc_is_i64", i32), ("num_tokens", i64), ("src_dtype", i32), ("dst_dtype", i32),
        ("k_scale", f32), ("v_scale", f32), ("size_limit", i64), ("skip_index", i64),
    ]


ROPE_FP32, ROPE_NATIVE = 0, 1
CACHE_F32, CACHE_MODEL_DTYPE = 0, 1


class DecodePlanArgs(C.Structure):
    _fields_ = [
        ("seq_lens", vp), ("seq_lens_is_i64", i32), ("batch_size", i32), ("max_bs", i32),
        ("max_chunks", i32), ("num_kv_heads", i32), ("num_ctas", i32), ("waves", i32),
        ("min_chunk_tokens", i32), ("fixed_chunk_tokens", i32), ("plan", vp), ("window_keys", i32),
    ]


class DraftTableArgs(C.Structure):
    _fields_ = [
        ("req_to_token", vp), ("req_to_token_stride", i64),
        ("req_pool_indices", vp), ("req_pool_indices_is_i64", i32),
        ("seq_lens", vp), ("seq_lens_is_i64", i32),
        ("num_seqs", i32), ("topk", i32), ("num_steps", i32), ("page_size", i32),
        ("table", vp), ("table_stride", i64),
    ]


class DecodeArgs(C.Structure):
    _fields_ = [
        ("q", vp), ("o", vp),
        ("q_stride_t", i64), ("q_stride_h", i64), ("o_stride_t", i64), ("o_stride_h", i64),
        ("k_cache", vp), ("v_cache", vp),
        ("k_stride_s", i64), ("k_stride_h", i64), ("v_stride_s", i64), ("v_stride_h", i64),
        ("req_to_token", vp), ("req_to_token_stride", i64),
        ("req_pool_indices", vp), ("req_pool_indices_is_i64", i32),
        ("seq_lens", vp), ("seq_lens_is_i64", i32),
        ("batch_size", i32), ("num_q_heads", i32), ("num_kv_heads", i32),
        ("head_dim_qk", i32), ("head_dim_v", i32), ("q_dtype", i32), ("kv_dtype", i32),
        ("sm_scale", f32), ("k_scale", f32), ("v_scale", f32), ("logit_cap", f32),
        ("plan", vp), ("max_bs", i32), ("max_chunks", i32),
        ("ws_o", vp), ("ws_lse", vp), ("num_ctas", i32), ("kernel", i32),
        ("ev_attn_start", vp), ("ev_attn_stop", vp),
        ("window_keys", i32), ("sinks", vp), ("lse", vp),
    ]


class ExtendPlanArgs(C.Structure):
    _fields_ = [
        ("extend_seq_lens", vp), ("lens_is_i64", i32), ("batch_size", i32), ("max_bs", i32),
        ("max_tiles", i32), ("block_q", i32), ("plan", vp),
        ("req_to_token", vp), ("req_to_token_stride", i64), ("req_pool_indices", vp),
        ("req_pool_indices_is_i64", i32), ("extend_prefix_lens", vp), ("page_size", i32),
    ]


class ExtendArgs(C.Structure):
    _fields_ = [
        ("q", vp), ("o", vp),
        ("q_stride_t", i64), ("q_stride_h", i64), ("o_stride_t", i64), ("o_stride_h", i64),
        ("k_cache", vp), ("v_cache", vp),
        ("k_stride_s", i64), ("k_stride_h", i64), ("v_stride_s", i64), ("v_stride_h", i64),
        ("req_to_token", vp), ("req_to_token_stride", i64),
        ("req_pool_indices", vp), ("req_pool_indices_is_i64", i32),
        ("extend_prefix_lens", vp), ("extend_seq_lens", vp), ("lens_is_i64", i32),
        ("batch_size", i32), ("num_q_tokens", i64),
        ("num_q_heads", i32), ("num_kv_heads", i32), ("head_dim_qk", i32), ("head_dim_v", i32),
        ("q_dtype", i32), ("kv_dtype", i32),
        ("sm_scale", f32), ("k_scale", f32), ("v_scale", f32), ("logit_cap", f32),
        ("causal", i32), ("plan", vp), ("max_bs", i32), ("max_tiles", i32),
        ("num_ctas", i32), ("kernel", i32), ("lse", vp), ("pool_slots", i64),
        ("window_keys", i32), ("sinks", vp), ("custom_mask", vp), ("mask_indptr", vp), ("mask_skip_prefix", i32),
        ("block_q", i32), ("page_size", i32),
        ("num_splits", i32), ("split_ws", vp), ("split_ws_floats", i64), ("num_tiles_hint", i32),
    ]


class MergeStateArgs(C.Structure):
    _fields_ = [
        ("o_a", vp), ("lse_a", vp), ("o_b", vp), ("lse_b", vp), ("o", vp), ("lse", vp),
        ("num_tokens", i64), ("num_heads", i32), ("head_dim", i32), ("dtype", i32),
        ("oa_stride_t", i64), ("oa_stride_h", i64), ("ob_stride_t", i64), ("ob_stride_h", i64),
        ("o_stride_t", i64), ("o_stride_h", i64),
    ]


MAX_WORLD = 8
OAR_AUTO, OAR_ONE_SHOT, OAR_TWO_SHOT = 0, 1, 2


class OProjAllReduceArgs(C.Structure):
    _fields_ = [
        ("x", vp), ("w", vp), ("x_stride", i64), ("w_stride", i64),
        ("num_tokens", i32), ("out_features", i32), ("in_features", i32), ("dtype", i32),
        ("rank", i32), ("world", i32), ("reserved0", C.c_uint32), ("max_tiles", i32),
        ("staging_peers", vp * MAX_WORLD), ("out_peers", vp * MAX_WORLD), ("flags_peers", vp * MAX_WORLD),
        ("staging_mc", vp), ("out_mc", vp), ("staging_bytes", i64), ("mode", i32), ("debug_times", vp),
    ]


# name -> (restype, argtypes); the test-suite checks this list against include/b200attn.h
SYMBOLS = {
    "b200attn_last_error": (C.c_char_p, []),
    "b200attn_abi_version": (C.c_int, []),
    "b200attn_device_info": (C.c_int, [C.POINTER(C.c_int)] * 3),
    "b200attn_store_kv": (C.c_int, [C.POINTER(StoreKVArgs), vp]),
    "b200attn_decode_plan_ints": (i64, [i32, i32]),
    "b200attn_draft_table": (C.c_int, [C.POINTER(DraftTableArgs), vp]),
    "b200attn_rope_store_kv": (C.c_int, [C.POINTER(RopeStoreArgs), vp]),
    "b200attn_decode_plan": (C.c_int, [C.POINTER(DecodePlanArgs), vp]),
    "b200attn_decode": (C.c_int, [C.POINTER(DecodeArgs), vp]),
    "b200attn_decode_static_ws": (C.c_int, [C.POINTER(DecodeArgs), vp]),
    "b200attn_decode_fast_supported": (C.c_int, [i32] * 5),
    "b200attn_decode_ctas_per_sm": (C.c_int, [i32, i32]),
    "b200attn_decode_items_per_chunk": (C.c_int, [i32] * 6),
    "b200attn_extend_plan_ints": (i64, [i32, i32]),
    "b200attn_extend_plan": (C.c_int, [C.POINTER(ExtendPlanArgs), vp]),
    "b200attn_extend": (C.c_int, [C.POINTER(ExtendArgs), vp]),
    "b200attn_extend_fast_supported": (C.c_int, [i32] * 5),
    "b200attn_extend_block_q": (C.c_int, [i32] * 6),
    "b200attn_extend_split_ws_floats": (i64, [i32, i64, i32, i32]),
    "b200attn_extend_pick_splits": (i32, [i32, i32, i32]),
    "b200attn_merge_state": (C.c_int, [C.POINTER(MergeStateArgs), vp]),
    "b200attn_oproj_allreduce": (C.c_int, [C.POINTER(OProjAllReduceArgs), vp]),
    "b200attn_oproj_allreduce_flag_words": (i64, [i32]),
    "b200attn_oproj_allreduce_tiles": (i32, [i32, i32]),
}


class B200AttnError(RuntimeError):
    pass


def _load() -> C.CDLL:
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m sglang_b200.build` "
            "(sglang_b200 has no CPU or PyTorch fallback)"
        )
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    got = lib.b200attn_abi_version()
    if got != ABI_VERSION:
        raise ImportError(f"{LIB_PATH}: ABI version {got}, expected {ABI_VERSION}")
    return lib


lib = _load()


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib.b200attn_last_error().decode(errors="replace")
        raise B200AttnError(f"{what} failed (rc={rc}): {msg}")
