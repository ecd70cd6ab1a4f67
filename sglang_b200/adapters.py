"""Callers for the two fused kernels that sit next to the attention path, with the reference's
own signatures, so a maintainer can bind them with one assignment each (INTEGRATION.md sections 6 and 7).

* `FusedSetKVBufferArg`, `apply_rope_with_cos_sin_cache_inplace(q, k, cos_sin_cache, positions, *, is_neox,
  rope_dim, fused_args)` -- reference `kernels/ops/attention/rope.py:97-115,183-227`; built per layer by
  `create_fused_set_kv_buffer_arg` (`srt/models/utils.py:309-330`) and passed down by
  `RotaryEmbedding.forward_cuda` (`srt/layers/rotary_embedding/base.py:363-385`).  One launch of
  `b200attn_rope_store_kv` instead of rotary + `set_kv_buffer`.
* `B200RowParallelLinear.forward(input_) -> (output, output_bias)` -- the o_proj call of the attention block,
  reference `RowParallelLinear.forward` (`srt/layers/linear.py:1563-1634`): GEMM on this rank's K shard followed by
  `tensor_model_parallel_all_reduce`.  Decode-sized inputs take ONE `b200attn_oproj_allreduce` launch; prefill-sized
  ones (T > `fused_max_tokens`) keep the library path (cuBLAS + NCCL), where the fused kernel measured 0.64-0.96x.

No CPU fallback: both raise on non-CUDA tensors like every op of this package.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from . import ops
from .comm import OProjAllReduce


@dataclass
class FusedSetKVBufferArg:
    """Field for field the reference's dataclass (`kernels/ops/attention/rope.py:97-115`)."""

    value: torch.Tensor      # (nnz, num_v_heads * head_size)
    k_buffer: torch.Tensor   # (slots, num_k_heads * head_size): the layer's K pool, flattened per row
    v_buffer: torch.Tensor   # (slots, num_v_heads * head_size)
    cache_loc: torch.Tensor  # (nnz,) pool rows = ForwardBatch.out_cache_loc


def create_fused_set_kv_buffer_arg(value: torch.Tensor, layer, forward_batch) -> FusedSetKVBufferArg:
    """Mirror of `srt/models/utils.py:309-330` (CUDA branch): the layer's pool buffers flattened per row."""
    pool = forward_batch.token_to_kv_pool
    if getattr(layer, "k_scale", None) is not None or getattr(layer, "v_scale", None) is not None:
        raise AssertionError("scale not supported")          # the reference asserts the same
    k_buffer = pool.get_key_buffer(layer.layer_id)
    v_buffer = pool.get_value_buffer(layer.layer_id)
    return FusedSetKVBufferArg(
        value=value,
        k_buffer=k_buffer.view(k_buffer.shape[0], -1),
        v_buffer=v_buffer.view(v_buffer.shape[0], -1),
        cache_loc=forward_batch.out_cache_loc,
    )


def apply_rope_with_cos_sin_cache_inplace(
    q: torch.Tensor,
    k: torch.Tensor,
    cos_sin_cache: torch.Tensor,
    positions: torch.Tensor,
    *,
    is_neox: bool,
    rope_dim: int = 0,
    fused_args: Optional[FusedSetKVBufferArg] = None,
) -> None:
    """RoPE on q [T, Hq, D] and k [T, Hkv, D] in place; with `fused_args` the rotated k and `fused_args.value` are
    also written to pool rows `cache_loc` in the same launch (same argument meaning as the reference function of
    this name).  `cos_sin_cache` [max_position, rope_dim], cos in the first half; float32 as the reference
    requires, or the model dtype."""
    if q.dim() != 3 or k.dim() != 3:
        raise ValueError("q and k must be [num_tokens, num_heads, head_dim]")
    head_dim = q.shape[-1]
    rope_dim = rope_dim or cos_sin_cache.size(-1)
    if rope_dim != cos_sin_cache.size(-1):
        raise ValueError(f"rope_dim {rope_dim} != cos_sin_cache width {cos_sin_cache.size(-1)}")
    if fused_args is None:
        ops.rope_store_kv(positions, q, k, None, cos_sin_cache, head_dim, is_neox=is_neox)
        return
    kb, vb = fused_args.k_buffer, fused_args.v_buffer
    T = q.shape[0]
    v = fused_args.value.view(T, k.shape[1], -1)
    ops.rope_store_kv(
        positions, q, k, v, cos_sin_cache, head_dim, is_neox=is_neox,
        k_cache=kb.view(kb.shape[0], k.shape[1], head_dim),
        v_cache=vb.view(vb.shape[0], k.shape[1], v.shape[-1]),
        loc=fused_args.cache_loc, cache_dtype=kb.dtype,
    )


class B200RowParallelLinear(torch.nn.Module):
    """`RowParallelLinear` for the attention block's o_proj with the all-reduce fused into the GEMM kernel.

    Same constructor vocabulary and the same `forward` contract as the reference class
    (`srt/layers/linear.py:1408-1634`): `input_` is this rank's `[T, input_size / tp_size]` slice when
    `input_is_parallel` (what attention hands over), the weight is `[output_size, input_size / tp_size]`, the result
    `(output, output_bias)` with `output` reduced over the TP group unless `reduce_results=False` or
    `skip_all_reduce=True`.  The bias is added by rank 0's share only when `skip_bias_add` is False -- here: after
    the reduction, which is arithmetically the same and keeps the kernel bias-free.
    """

    def __init__(self, input_size: int, output_size: int, bias: bool = False, input_is_parallel: bool = True,
                 skip_bias_add: bool = False, params_dtype: Optional[torch.dtype] = None,
                 reduce_results: bool = True, tp_rank: Optional[int] = None, tp_size: Optional[int] = None,
                 group=None, max_tokens: int = 8192, fused_max_tokens: int = 256, device="cuda"):
        super().__init__()
        import torch.distributed as dist

        if not input_is_parallel:
            raise NotImplementedError("the attention o_proj always receives its own K shard (input_is_parallel)")
        if tp_size is None:
            tp_size = dist.get_world_size(group) if dist.is_initialized() else 1
        if tp_rank is None:
            tp_rank = dist.get_rank(group) if dist.is_initialized() else 0
        if input_size % tp_size:
            raise ValueError(f"input_size {input_size} is not divisible by tp_size {tp_size}")
        self.input_size, self.output_size = input_size, output_size
        self.input_size_per_partition = input_size // tp_size
        self.tp_rank, self.tp_size, self.group = tp_rank, tp_size, group
        self.skip_bias_add, self.reduce_results = skip_bias_add, reduce_results
        dtype = params_dtype or torch.bfloat16
        self.weight = torch.nn.Parameter(
            torch.empty(output_size, self.input_size_per_partition, dtype=dtype, device=device), requires_grad=False)
        self.bias = torch.nn.Parameter(torch.zeros(output_size, dtype=dtype, device=device),
                                       requires_grad=False) if bias else None
        #: largest T that takes the fused kernel; above it cuBLAS + NCCL is faster
        self.fused_max_tokens = int(fused_max_tokens)
        self._comm = OProjAllReduce(group if tp_size > 1 else None, min(max_tokens, self.fused_max_tokens), output_size,
                                    dtype, device)

    def forward(self, input_: torch.Tensor, skip_all_reduce: bool = False, forward_batch=None,
                output_tensor: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        import torch.distributed as dist

        x = input_.reshape(-1, self.input_size_per_partition)
        reduce = self.reduce_results and self.tp_size > 1 and not skip_all_reduce
        bias_ = None if (self.tp_rank > 0 or self.skip_bias_add) else self.bias
        if x.shape[0] <= self.fused_max_tokens and (reduce or self.tp_size == 1):
            out = self._comm.forward(x, self.weight)          # GEMM + all-reduce, one launch
            out = out.clone() if output_tensor is None else output_tensor.copy_(out)
            if self.bias is not None and not self.skip_bias_add:
                out += self.bias                               # once, after the sum (rank 0's share in the reference)
        else:
            out = torch.nn.functional.linear(x, self.weight, bias_)
            if output_tensor is not None:
                out = output_tensor.copy_(out)
            if reduce:
                dist.all_reduce(out, group=self.group)
        out = out.view(*input_.shape[:-1], self.output_size)
        return out, (self.bias if self.skip_bias_add else None)
