"""Tensor-parallel head partitioning for the attention path (host logic only).

SGLang shards Q heads and KV heads across TP ranks and replicates a KV head when
tp_size > num_kv_heads (reference srt/models/llama.py:158-171, srt/configs/model_config.py:1121-1128).
Every rank's pool holds only its KV-head slice of every token; the page table and out_cache_loc
are identical on all ranks, so store/decode/extend need no communication.  The one collective of
the block is the o_proj all-reduce after attention (srt/layers/linear.py:1563), outside this path.
"""
from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class HeadShard:
    q_heads: int          # query heads on this rank
    kv_heads: int         # kv heads on this rank (>= 1)
    q_head_start: int     # first global query head
    kv_head_start: int    # first global kv head
    kv_replicas: int      # ranks holding the same kv head (tp_size > total kv heads)


def shard_heads(total_q_heads: int, total_kv_heads: int, tp_size: int, tp_rank: int) -> HeadShard:
    if tp_size <= 0 or not 0 <= tp_rank < tp_size:
        raise ValueError(f"bad tp rank {tp_rank}/{tp_size}")
    if total_q_heads % tp_size != 0:
        raise ValueError(f"{total_q_heads} query heads are not divisible by tp_size {tp_size}")
    q = total_q_heads // tp_size
    if total_kv_heads >= tp_size:
        if total_kv_heads % tp_size != 0:
            raise ValueError(f"{total_kv_heads} kv heads are not divisible by tp_size {tp_size}")
        kv = total_kv_heads // tp_size
        return HeadShard(q, kv, tp_rank * q, tp_rank * kv, 1)
    if tp_size % total_kv_heads != 0:
        raise ValueError(f"tp_size {tp_size} is not a multiple of {total_kv_heads} kv heads")
    rep = tp_size // total_kv_heads
    return HeadShard(q, 1, tp_rank * q, tp_rank // rep, rep)
