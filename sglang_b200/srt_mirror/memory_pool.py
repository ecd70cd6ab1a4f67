"""ReqToTokenPool / MHATokenToKVPool / KVWriteLoc: the page table and the token-granular
KV pool (reference srt/mem_cache/memory_pool.py:256-283, 1532-1571, 1740-2430).

Same buffer layout as the reference's NHD pool: per layer K `[size + page_size, Hkv, Dk]`
and V `[size + page_size, Hkv, Dv]`, FP8 stored as uint8 (`store_dtype`), slot 0 / page 0
reserved as the sink that padded CUDA-graph rows write to.  `set_kv_buffer` launches this
repo's scatter kernel; there is no `k_cache[loc] = k` fallback.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch

_FP8 = (torch.float8_e4m3fn, torch.float8_e5m2)


class ReqToTokenPool:
    def __init__(self, size: int, max_context_len: int, device: str, enable_memory_saver: bool = False):
        self.size = size
        self._alloc_size = size + 1  # row 0 is the dummy row padded graph batches point at
        self.max_context_len = max_context_len
        self.device = device
        self.req_to_token = torch.zeros((self._alloc_size, max_context_len), dtype=torch.int32, device=device)
        self.free_slots = list(range(1, self._alloc_size))

    def write(self, indices, values):
        self.req_to_token[indices] = values

    def available_size(self) -> int:
        return len(self.free_slots)

    def alloc(self, need_size: int) -> Optional[List[int]]:
        if need_size > len(self.free_slots):
            return None
        out, self.free_slots = self.free_slots[:need_size], self.free_slots[need_size:]
        return out

    def free(self, idx) -> None:
        self.free_slots.extend([idx] if isinstance(idx, int) else list(idx))

    def clear(self) -> None:
        self.free_slots = list(range(1, self._alloc_size))


@dataclass
class KVWriteLoc:
    loc: torch.Tensor
    swa_loc: Optional[torch.Tensor] = None
    full_loc: Optional[torch.Tensor] = None


def unwrap_write_loc(loc_info):
    if isinstance(loc_info, KVWriteLoc):
        return loc_info.loc, loc_info.swa_loc, loc_info.full_loc
    return loc_info, None, None


class MHATokenToKVPool:
    def __init__(self, size: int, page_size: int, dtype: torch.dtype, head_num: int, head_dim: int,
                 layer_num: int, device: str, enable_memory_saver: bool = False,
                 v_head_dim: Optional[int] = None, start_layer: Optional[int] = None,
                 end_layer: Optional[int] = None, enable_alt_stream: bool = False):
        self.size = size
        self.page_size = page_size
        self.dtype = dtype
        self.store_dtype = torch.uint8 if dtype in _FP8 else dtype
        self.device = device
        self.head_num = head_num
        self.head_dim = head_dim
        self.v_head_dim = v_head_dim if v_head_dim is not None else head_dim
        self.layer_num = layer_num
        self.start_layer = start_layer or 0
        self.end_layer = end_layer or layer_num - 1
        self.layer_transfer_counter = None
        rows = size + page_size
        self.k_buffer = [torch.zeros((rows, head_num, head_dim), dtype=self.store_dtype, device=device)
                         for _ in range(layer_num)]
        self.v_buffer = [torch.zeros((rows, head_num, self.v_head_dim), dtype=self.store_dtype, device=device)
                         for _ in range(layer_num)]

    # ---- reads (memory_pool.py:2257-2303)
    def _view(self, buf: torch.Tensor) -> torch.Tensor:
        return buf.view(self.dtype) if self.store_dtype != self.dtype else buf

    def get_key_buffer(self, layer_id: int) -> torch.Tensor:
        return self._view(self.k_buffer[layer_id - self.start_layer])

    def get_value_buffer(self, layer_id: int) -> torch.Tensor:
        return self._view(self.v_buffer[layer_id - self.start_layer])

    def get_kv_buffer(self, layer_id: int):
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def get_kv_size_bytes(self):
        k = sum(b.numel() * b.element_size() for b in self.k_buffer)
        v = sum(b.numel() * b.element_size() for b in self.v_buffer)
        return k, v

    # ---- write (memory_pool.py:2305-2381)
    def set_kv_buffer(self, layer, loc_info, cache_k: torch.Tensor, cache_v: torch.Tensor,
                      k_scale: Optional[float] = None, v_scale: Optional[float] = None,
                      layer_id_override: Optional[int] = None):
        from .. import ops

        loc, _, _ = unwrap_write_loc(loc_info)
        layer_id = layer_id_override if layer_id_override is not None else layer.layer_id
        idx = layer_id - self.start_layer
        same = cache_k.dtype == self.dtype
        ops.store_kv(
            cache_k, cache_v, self.k_buffer[idx], self.v_buffer[idx], loc,
            cache_dtype=self.dtype,
            k_scale=None if same else k_scale, v_scale=None if same else v_scale,
            size_limit=self.size + self.page_size, skip_index=0,
        )
