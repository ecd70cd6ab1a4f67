"""Tensor-parallel o_proj with the all-reduce fused into the GEMM kernel.

Host-side mirror of the reference's `RowParallelLinear.forward` -> `tensor_model_parallel_all_reduce`
(srt/layers/linear.py:1563, srt/distributed/communication_op.py:18-20, parallel_state.py:622) for the
attention block's output projection.  The device work is ONE hand-written sm_100a kernel per rank
(`b200attn_oproj_allreduce`, csrc/oproj_allreduce.cu): tcgen05 GEMM tiles whose epilogue exchanges and
reduces the partial products through NVLink / NVSwitch peer memory (multimem.ld_reduce / multimem.st when
a multicast mapping exists).  PyTorch is plumbing only: `torch.distributed._symmetric_memory` allocates the
symmetric buffer and exchanges the peer / multicast addresses; no NCCL collective runs on the data path.

There is no fallback: without the CUDA library, without a process group, or on CPU tensors this raises.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from . import _C, ops


def tile_owner(tile: int, world: int) -> int:
    """Rank that reduces output tile `tile` (tiles are numbered row-major over [ceil(T/128), N/64])."""
    return tile % world


def num_tiles(num_tokens: int, out_features: int) -> int:
    return ((num_tokens + 127) // 128) * ((out_features + 63) // 64)


def symmetric_layout(max_tokens: int, out_features: int, itemsize: int, max_tiles: int):
    """Byte offsets of the regions inside the one symmetric allocation: staging x 2 | out | flags."""
    region = (max_tokens * out_features * itemsize + 1023) // 1024 * 1024
    flag_bytes = (int(_C.lib.b200attn_oproj_allreduce_flag_words(max_tiles)) * 4 + 1023) // 1024 * 1024
    return {"staging": 0, "staging_bytes": region, "out": 2 * region, "flags": 3 * region,
            "total": 3 * region + flag_bytes}


class OProjAllReduce:
    """`out = all_reduce(x @ w.T)` for one TP group; owns the symmetric staging / output / flag buffers.

    `forward` returns a view of the symmetric output buffer (valid until the next call), exactly like the
    reference's custom all-reduce returns its registered buffer."""

    def __init__(self, group, max_tokens: int, out_features: int, dtype: torch.dtype, device,
                 use_multicast: bool = True, mode: int = _C.OAR_AUTO):
        import torch.distributed as dist

        self.group = group
        if group is None and not dist.is_initialized():      # single GPU: the GEMM alone (TP 1)
            self.rank, self.world = 0, 1
        else:
            self.rank = dist.get_rank(group)
            self.world = dist.get_world_size(group)
        if self.world > _C.MAX_WORLD:
            raise ValueError(f"TP world size {self.world} > {_C.MAX_WORLD}")
        if out_features % 64:
            raise ValueError("out_features must be a multiple of 64")
        self.max_tokens, self.N, self.dtype = int(max_tokens), int(out_features), dtype
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("OProjAllReduce needs a CUDA device (sm_100a); there is no CPU fallback")
        self.max_tiles = num_tiles(self.max_tokens, self.N)
        self.lay = symmetric_layout(self.max_tokens, self.N, dtype.itemsize, self.max_tiles)
        self.mode = mode
        self._mc = 0
        #: optional int64 [max_tiles, 8] tensor the kernel fills with %globaltimer stamps (tuning aid)
        self.debug_times: Optional[torch.Tensor] = None
        if self.world == 1:
            self._buf = torch.zeros(self.lay["total"], dtype=torch.uint8, device=self.device)
            self._peer_ptrs: List[int] = [self._buf.data_ptr()]
        else:
            import torch.distributed._symmetric_memory as symm

            self._buf = symm.empty(self.lay["total"], dtype=torch.uint8, device=self.device)
            self._hdl = symm.rendezvous(self._buf, group)
            self._peer_ptrs = [int(p) for p in self._hdl.buffer_ptrs]
            if use_multicast:
                self._mc = int(getattr(self._hdl, "multicast_ptr", 0) or 0)
            self._buf.zero_()                       # flags and the device-side launch counter start at 0
            torch.cuda.synchronize(self.device)
            dist.barrier(group)
        lo = self.lay["out"]
        n_out = self.max_tokens * self.N * dtype.itemsize
        self._out = self._buf[lo : lo + n_out].view(dtype).view(self.max_tokens, self.N)

    @property
    def multicast(self) -> bool:
        return self._mc != 0

    def forward(self, x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
        """x [T, K_shard], w [N, K_shard] (torch Linear layout) -> reduced [T, N] view."""
        if not (x.is_cuda and w.is_cuda):
            raise RuntimeError("sglang_b200 ops run only on CUDA tensors (no CPU fallback)")
        if x.dim() != 2 or w.dim() != 2 or x.shape[1] != w.shape[1] or w.shape[0] != self.N:
            raise ValueError(f"shapes: x {tuple(x.shape)}, w {tuple(w.shape)}, N {self.N}")
        if x.dtype != self.dtype or w.dtype != self.dtype or x.stride(1) != 1 or w.stride(1) != 1:
            raise ValueError("x and w must have the communicator's dtype and unit inner stride")
        T, K = x.shape
        if T > self.max_tokens:
            raise ValueError(f"{T} tokens exceed the communicator's capacity {self.max_tokens}")
        if T == 0:
            return self._out[:0]
        a = _C.OProjAllReduceArgs(
            x=x.data_ptr(), w=w.data_ptr(), x_stride=x.stride(0), w_stride=w.stride(0),
            num_tokens=T, out_features=self.N, in_features=K, dtype=ops.dtype_code(self.dtype),
            rank=self.rank, world=self.world, max_tiles=self.max_tiles,
            staging_mc=(self._mc + self.lay["staging"]) if self._mc else None,
            out_mc=(self._mc + self.lay["out"]) if self._mc else None,
            staging_bytes=self.lay["staging_bytes"], mode=self.mode,
            debug_times=self.debug_times.data_ptr() if self.debug_times is not None else None,
        )
        for r, base in enumerate(self._peer_ptrs):
            a.staging_peers[r] = base + self.lay["staging"]
            a.out_peers[r] = base + self.lay["out"]
            a.flags_peers[r] = base + self.lay["flags"]
        _C.check(_C.lib.b200attn_oproj_allreduce(C.byref(a), torch.cuda.current_stream().cuda_stream),
                 "oproj_allreduce")
        ops.launch_count += 1
        return self._out[:T]


def oproj_allreduce_ref(xs, ws, with_scale: bool = False):
    """What the fused kernel must equal (used by tests/benchmarks): every rank's partial product rounded to the
    model dtype -- the RowParallelLinear output the reference all-reduces -- summed in float32, rounded once.
    `with_scale` also returns sum_r |partial_r|: a partial that sits on a rounding boundary may land one ulp
    either way (fp32 vs float64 accumulation), so the comparison tolerance is an ulp of that magnitude."""
    dtype = xs[0].dtype
    acc: Optional[torch.Tensor] = None
    mag: Optional[torch.Tensor] = None
    for x, w in zip(xs, ws):
        part = (x.double() @ w.double().t()).to(torch.float32).to(dtype).to(torch.float32)
        acc = part if acc is None else acc + part
        mag = part.abs() if mag is None else mag + part.abs()
    return (acc.to(dtype), mag) if with_scale else acc.to(dtype)
