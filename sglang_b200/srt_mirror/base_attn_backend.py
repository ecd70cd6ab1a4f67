"""AttentionBackend plugin API (reference srt/layers/attention/base_attn_backend.py:21-281):
the three-method metadata contract, the CUDA-graph hooks and the mode dispatch in `forward`."""
from __future__ import annotations

from abc import ABC


class AttentionBackend(ABC):
    needs_cpu_seq_lens: bool = True
    # class-level switches the runners read (base_attn_backend.py:45-48,102-115)
    supports_ragged_verify_graph: bool = False
    use_captured_forward_metadata_for_breakable_cuda_graph: bool = False
    supports_full_cuda_graph_chunked_prefix: bool = False

    def draft_extend_metadata_captured_in_graph(self) -> bool:
        return False

    @property
    def verify_mask(self):
        """The mask the draft stage fills in place, if the backend owns one (:172-175)."""
        return None

    def update_verify_buffers_to_fill_after_draft(self, spec_info, cuda_graph_bs):
        raise NotImplementedError()

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
