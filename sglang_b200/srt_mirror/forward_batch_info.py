"""ForwardMode / ForwardBatch: the fields of the reference batch object that attention
backends read (reference srt/model_executor/forward_batch_info.py:98-197, 412-553)."""
from __future__ import annotations

from dataclasses import dataclass
from enum import IntEnum, auto
from typing import List, Optional

import torch


class ForwardMode(IntEnum):
    EXTEND = auto()
    DECODE = auto()
    MIXED = auto()
    IDLE = auto()
    TARGET_VERIFY = auto()
    DRAFT_EXTEND_V2 = auto()
    PREBUILT = auto()
    SPLIT_PREFILL = auto()
    DLLM_EXTEND = auto()

    def is_extend(self, include_draft_extend_v2: bool = False) -> bool:
        return self in (ForwardMode.EXTEND, ForwardMode.MIXED, ForwardMode.TARGET_VERIFY,
                        ForwardMode.SPLIT_PREFILL, ForwardMode.DLLM_EXTEND) or (
            include_draft_extend_v2 and self == ForwardMode.DRAFT_EXTEND_V2)

    def is_prefill(self, include_draft_extend_v2: bool = False) -> bool:
        return self.is_extend(include_draft_extend_v2)

    def is_decode(self) -> bool:
        return self == ForwardMode.DECODE

    def is_mixed(self) -> bool:
        return self == ForwardMode.MIXED

    def is_idle(self) -> bool:
        return self == ForwardMode.IDLE

    def is_decode_or_idle(self) -> bool:
        return self in (ForwardMode.DECODE, ForwardMode.IDLE)

    def is_target_verify(self) -> bool:
        return self == ForwardMode.TARGET_VERIFY

    def is_draft_extend_v2(self) -> bool:
        return self == ForwardMode.DRAFT_EXTEND_V2

    def is_cuda_graph(self) -> bool:
        return self in (ForwardMode.DECODE, ForwardMode.TARGET_VERIFY, ForwardMode.IDLE, ForwardMode.DLLM_EXTEND)


@dataclass
class ForwardBatch:
    forward_mode: ForwardMode
    batch_size: int
    req_pool_indices: torch.Tensor            # [bs] int32/int64 rows of req_to_token
    seq_lens: torch.Tensor                    # [bs] total KV length per request (prefix + new)
    out_cache_loc: torch.Tensor               # [num new tokens] slot ids to write
    seq_lens_sum: int = 0
    seq_lens_cpu: Optional[torch.Tensor] = None
    input_ids: Optional[torch.Tensor] = None
    positions: Optional[torch.Tensor] = None
    # extend only
    extend_prefix_lens: Optional[torch.Tensor] = None
    extend_prefix_lens_cpu: Optional[List[int]] = None
    extend_seq_lens: Optional[torch.Tensor] = None
    extend_seq_lens_cpu: Optional[List[int]] = None
    extend_num_tokens: Optional[int] = None
    # not used by dense MHA configs; kept so backends can test them for None
    encoder_lens: Optional[torch.Tensor] = None
    encoder_out_cache_loc: Optional[torch.Tensor] = None
    spec_info: Optional[object] = None       # TARGET_VERIFY: .draft_token_num, .custom_mask
    mha_return_lse: bool = False             # extend returns (o, lse) (radix_attention.py:244-277)
    _attn_output: Optional[torch.Tensor] = None
