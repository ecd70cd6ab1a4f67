"""AttentionBackend plugin API (reference srt/layers/attention/base_attn_backend.py:21-281):
the three-method metadata contract, the CUDA-graph hooks and the mode dispatch in `forward`."""
from __future__ import annotations

from abc import ABC


class AttentionBackend(ABC):
    needs_cpu_seq_lens: bool = True

    def init_forward_metadata(self, forward_batch):
        """Eager entry point: `_out_graph` then `_in_graph` (base_attn_backend.py:50-56)."""
        self.init_forward_metadata_out_graph(forward_batch)
        self.init_forward_metadata_in_graph(forward_batch)

    def init_forward_metadata_out_graph(self, forward_batch, in_capture: bool = False):
        """Host-side per-iteration prep, outside graph capture (:58-81)."""

    def init_forward_metadata_in_graph(self, forward_batch):
        """Graph-recordable static-shape device work (:83-95)."""

    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int):
        raise NotImplementedError()

    def get_cuda_graph_seq_len_fill_value(self):
        raise NotImplementedError()

    def on_after_cuda_graph_warmup(self):
        pass

    def forward(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        """Mode dispatch (:189-231)."""
        mode = forward_batch.forward_mode
        if mode.is_idle():
            return q.new_empty(q.shape[0], layer.tp_q_head_num * layer.v_head_dim)
        if mode.is_decode():
            return self.forward_decode(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)
        return self.forward_extend(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)

    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        raise NotImplementedError()

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        raise NotImplementedError()

    def support_triton(self):
        return True
