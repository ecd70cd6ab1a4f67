"""The slice of `ModelRunner` an attention backend reads at construction (reference
srt/model_executor/model_runner.py; the unit-test kit's MockModelRunner,
test/kits/attention_unittest/attention_methods/dense_attention.py:306-418): device, dtype, page size, the two
pools, `model_config` head counts and context length, `sliding_window_size`, `server_args`."""
from __future__ import annotations

from types import SimpleNamespace
from typing import Optional

import torch

from .memory_pool import MHATokenToKVPool, ReqToTokenPool


def make_model_runner(*, device, dtype: torch.dtype, pool_dtype: torch.dtype, num_heads: int, num_kv_heads: int,
                      head_dim: int, page_size: int, pool_rows: int, max_bs: int, max_context_len: int,
                      layer_num: int = 1, v_head_dim: Optional[int] = None,
                      sliding_window_size: Optional[int] = None, speculative_num_draft_tokens: int = 0):
    """`pool_rows` counts the sink page: the pool gets `size = pool_rows - page_size` usable slots."""
    rtp = ReqToTokenPool(size=max_bs, max_context_len=max_context_len, device=device)
    pool = MHATokenToKVPool(size=pool_rows - page_size, page_size=page_size, dtype=pool_dtype, head_num=num_kv_heads,
                            head_dim=head_dim, layer_num=layer_num, device=device, v_head_dim=v_head_dim)
    cfg = SimpleNamespace(
        num_attention_heads=num_heads, num_key_value_heads=num_kv_heads, head_dim=head_dim,
        context_len=max_context_len,
        get_num_kv_heads=lambda tp: num_kv_heads // tp, get_max_num_attention_heads=lambda: num_heads,
    )
    return SimpleNamespace(device=device, dtype=dtype, page_size=page_size, model_config=cfg,
                           req_to_token_pool=rtp, token_to_kv_pool=pool, sliding_window_size=sliding_window_size,
                           server_args=SimpleNamespace(speculative_num_draft_tokens=speculative_num_draft_tokens))
