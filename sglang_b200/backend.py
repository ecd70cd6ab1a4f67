"""B200AttnBackend: the `--attention-backend b200` plugin.

Implements the reference AttentionBackend API (srt/layers/attention/base_attn_backend.py:21-281)
over the unchanged ForwardBatch, ReqToTokenPool page table and MHATokenToKVPool:

    init_forward_metadata{,_out_graph,_in_graph}   one planner launch per forward, on the device
    forward_decode                                 KV-store scatter + split-KV decode + merge
    forward_extend                                 KV-store scatter + prefix/causal extend
    init_cuda_graph_state / get_cuda_graph_seq_len_fill_value   CUDA-graph contract

Behavioural model: torch_native_backend.py:279-402 (what is computed) and
triton_backend.py:572-634, 962-1063, 1250-1437, 1714-1864 (metadata + graph hooks).
All device work is hand-written sm_100a CUDA behind include/b200attn.h; there is no Triton,
no torch fallback, and construction fails on a non-CUDA device.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops

try:  # inside an SGLang process: subclass the real ABC so isinstance checks hold
    from sglang.srt.layers.attention.base_attn_backend import AttentionBackend as _Base  # type: ignore
except Exception:  # standalone (tests, bench, GPU box): the mirror of the same API
    from .srt_mirror.base_attn_backend import AttentionBackend as _Base


def _attn_tp_size(model_runner) -> int:
    try:
        from sglang.srt.runtime_context import get_parallel  # type: ignore

        return int(get_parallel().attn_tp_size)
    except Exception:
        return int(getattr(model_runner, "attn_tp_size", 1) or 1)


class B200AttnBackend(_Base):
    #: rows of the fast decode kernel's work list per persistent CTA
    DECODE_WAVES = 16

    def __init__(self, model_runner, skip_prefill: bool = False):
        super().__init__()
        dev = torch.device(model_runner.device)
        if dev.type != "cuda":
            raise RuntimeError("B200AttnBackend needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = dev
        sm_count, _, _ = ops.device_info(dev)
        self.sm_count = sm_count
        # pools are resolved through the backend by other code (forward_context.py:70-76)
        self.req_to_token_pool = model_runner.req_to_token_pool
        self.token_to_kv_pool = model_runner.token_to_kv_pool
        self.page_size = int(getattr(model_runner, "page_size", 1) or 1)
        self.skip_prefill = skip_prefill

        cfg = model_runner.model_config
        tp = _attn_tp_size(model_runner)
        if hasattr(cfg, "get_max_num_attention_heads"):
            self.num_head = cfg.get_max_num_attention_heads() // tp
        else:
            self.num_head = cfg.num_attention_heads // tp
        self.num_kv_head = cfg.get_num_kv_heads(tp)
        pool = self.token_to_kv_pool
        if not (isinstance(getattr(pool, "k_buffer", None), list) and pool.k_buffer[0].dim() == 3):
            raise NotImplementedError(
                "b200 backend supports the NHD MHATokenToKVPool layout [slots, Hkv, D] only")
        self.head_dim = pool.k_buffer[0].shape[-1]
        self.v_head_dim = pool.v_buffer[0].shape[-1]
        self.kv_dtype = pool.dtype
        self._q_dtype = getattr(model_runner, "dtype", torch.bfloat16)
        if getattr(model_runner, "sliding_window_size", None) not in (None, -1):
            raise NotImplementedError("b200 backend: sliding-window layers are not supported yet")

        self.max_bs = int(self.req_to_token_pool.size)
        self.max_context_len = int(getattr(cfg, "context_len", self.req_to_token_pool.max_context_len))
        self.decode_kernel = ops.KERNEL_AUTO
        self.extend_kernel = ops.KERNEL_AUTO
        self.decode_ctas = sm_count * ops.decode_ctas_per_sm(self.head_dim, self.kv_dtype)
        self._decode_items = ops.decode_items_per_chunk(self.num_head, self.num_kv_head, self.head_dim,
                                                        self.v_head_dim, self._q_dtype, self.kv_dtype)
        self.decode_ws = ops.DecodeWorkspace.create(
            self.max_bs, self.num_head, self.v_head_dim, dev, num_ctas=self.decode_ctas, waves=self.DECODE_WAVES)
        self.extend_ws: Optional[ops.ExtendWorkspace] = None
        self._extend_block_q = 64
        self.forward_metadata = None
        #: optional (start, stop) torch.cuda.Event pair re-recorded around the next decode attention kernel
        self.attn_events = None

    # ------------------------------------------------------------------ metadata
    def init_forward_metadata_out_graph(self, forward_batch, in_capture: bool = False):
        """Host-side prep only: (re)size the extend work list from host-known counts.

        Decode needs nothing here: the planner kernel reads `seq_lens` on the device, so a
        graph replay only has to refresh the runner's static `seq_lens`/`req_pool_indices`."""
        mode = forward_batch.forward_mode
        if mode.is_decode_or_idle():
            if forward_batch.batch_size > self.max_bs:
                raise ValueError(f"batch size {forward_batch.batch_size} exceeds pool size {self.max_bs}")
            return
        T = int(forward_batch.extend_num_tokens if forward_batch.extend_num_tokens is not None
                else forward_batch.out_cache_loc.shape[0])
        bs = int(forward_batch.batch_size)
        group = self.num_head // self.num_kv_head
        q_dtype = self._q_dtype_hint
        self._extend_block_q = ops.extend_block_q(self.head_dim, self.v_head_dim, group, q_dtype,
                                                  self.kv_dtype, self.extend_kernel)
        need_tiles = T // self._extend_block_q + bs + 1
        ws = self.extend_ws
        if ws is None or ws.max_tiles < need_tiles or ws.max_bs < bs:
            self.extend_ws = ops.ExtendWorkspace.create(
                max(self.max_bs, bs), max(need_tiles, 2 * (ws.max_tiles if ws else 0)), self.device,
                num_ctas=self.sm_count)

    def init_forward_metadata_in_graph(self, forward_batch):
        """One planner launch (static shape, graph-recordable, no host sync)."""
        mode = forward_batch.forward_mode
        if mode.is_idle():
            return
        if mode.is_decode():
            ops.decode_plan(self.decode_ws, forward_batch.seq_lens[: forward_batch.batch_size],
                            self.num_kv_head, waves=self.DECODE_WAVES, items_per_chunk=self._decode_items)
        else:
            ops.extend_plan(self.extend_ws, forward_batch.extend_seq_lens, self._extend_block_q)

    # model dtype is not on the runner contract; bf16 unless the runner says otherwise
    @property
    def _q_dtype_hint(self) -> torch.dtype:
        return getattr(self, "_q_dtype", torch.bfloat16)

    # ------------------------------------------------------------------ CUDA graphs
    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int):
        if max_bs > self.decode_ws.max_bs:
            self.max_bs = max_bs
            self._decode_items = ops.decode_items_per_chunk(self.num_head, self.num_kv_head, self.head_dim,
                                                        self.v_head_dim, self._q_dtype, self.kv_dtype)
        self.decode_ws = ops.DecodeWorkspace.create(
                max_bs, self.num_head, self.v_head_dim, self.device, num_ctas=self.decode_ctas,
                waves=self.DECODE_WAVES)

    def get_cuda_graph_seq_len_fill_value(self):
        return 1

    # ------------------------------------------------------------------ KV write
    def _save_kv(self, layer, cache_loc, k, v):
        """Scatter the new rows straight into the pool's layer buffers (same layout and
        cast rules as MHATokenToKVPool.set_kv_buffer, memory_pool.py:2305-2381; slot 0 is
        the sink for padded rows, kvcache.cuh:215)."""
        pool = self.token_to_kv_pool
        idx = layer.layer_id - getattr(pool, "start_layer", 0)
        same = k.dtype == pool.dtype
        ops.store_kv(
            k, v, pool.k_buffer[idx], pool.v_buffer[idx], cache_loc,
            cache_dtype=pool.dtype,
            k_scale=None if same else layer.k_scale_float,
            v_scale=None if same else layer.v_scale_float,
            size_limit=pool.size + pool.page_size, skip_index=0,
        )

    @staticmethod
    def _check_layer(layer):
        if getattr(layer, "is_cross_attention", False):
            raise NotImplementedError("b200 backend: cross attention is not supported")
        sw = getattr(layer, "sliding_window_size", -1)
        if sw is not None and sw > -1:
            raise NotImplementedError("b200 backend: sliding-window layers are not supported yet")

    def _descale(self, layer):
        if getattr(layer, "k_scale", None) is not None and getattr(layer, "v_scale", None) is not None:
            return float(layer.k_scale_float), float(layer.v_scale_float)
        return 1.0, 1.0

    def _pool_buffers(self, layer):
        pool = self.token_to_kv_pool
        # fetched per call: the pool may re-create its buffers (memory saver / resume)
        return pool.get_key_buffer(layer.layer_id), pool.get_value_buffer(layer.layer_id)

    # ------------------------------------------------------------------ forward
    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        self._check_layer(layer)
        if kwargs.get("sinks") is not None:
            raise NotImplementedError("b200 backend: attention sinks are not supported")
        q = q.reshape(-1, layer.tp_q_head_num * layer.qk_head_dim)
        if layer.qk_head_dim != layer.v_head_dim:
            o = q.new_empty((q.shape[0], layer.tp_q_head_num * layer.v_head_dim))
        else:
            o = torch.empty_like(q)
        if save_kv_cache and k is not None and v is not None:
            self._save_kv(layer, forward_batch.out_cache_loc, k, v)
        k_buf, v_buf = self._pool_buffers(layer)
        k_descale, v_descale = self._descale(layer)
        bs = q.shape[0]
        ops.decode_attention(
            q.view(bs, layer.tp_q_head_num, layer.qk_head_dim),
            o.view(bs, layer.tp_q_head_num, layer.v_head_dim),
            k_buf, v_buf,
            self.req_to_token_pool.req_to_token,
            forward_batch.req_pool_indices[:bs], forward_batch.seq_lens[:bs],
            self.decode_ws, layer.scaling,
            kv_dtype=self.kv_dtype, k_scale=k_descale, v_scale=v_descale,
            logit_cap=float(getattr(layer, "logit_cap", 0.0) or 0.0), kernel=self.decode_kernel,
            attn_events=self.attn_events,
        )
        return o

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        self._check_layer(layer)
        if kwargs.get("sinks") is not None:
            raise NotImplementedError("b200 backend: attention sinks are not supported")
        attn_out = getattr(forward_batch, "_attn_output", None)
        if attn_out is not None:
            o = attn_out
        elif layer.qk_head_dim != layer.v_head_dim:
            o = q.new_empty((q.shape[0], layer.tp_q_head_num * layer.v_head_dim))
        else:
            o = torch.empty_like(q)
        if save_kv_cache and k is not None and v is not None:
            self._save_kv(layer, forward_batch.out_cache_loc, k, v)
        k_buf, v_buf = self._pool_buffers(layer)
        k_descale, v_descale = self._descale(layer)
        from .srt_mirror.radix_attention import AttentionType as _AT

        attn_type = getattr(layer, "attn_type", None)
        causal = not (getattr(attn_type, "value", attn_type) == _AT.ENCODER_ONLY.value)
        T = q.shape[0]
        ops.extend_attention(
            q.view(T, layer.tp_q_head_num, layer.qk_head_dim),
            o.view(T, layer.tp_q_head_num, layer.v_head_dim),
            k_buf, v_buf,
            self.req_to_token_pool.req_to_token,
            forward_batch.req_pool_indices, forward_batch.extend_prefix_lens, forward_batch.extend_seq_lens,
            self.extend_ws, layer.scaling,
            kv_dtype=self.kv_dtype, k_scale=k_descale, v_scale=v_descale,
            logit_cap=float(getattr(layer, "logit_cap", 0.0) or 0.0), causal=causal,
            kernel=self.extend_kernel,
        )
        return o

    def support_triton(self):
        return True
