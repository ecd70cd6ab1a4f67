"""Host-side mirror of the reference interfaces the attention hot path touches.

The GPU box has no SGLang checkout, so the parity tests, `smoke()` and `bench.py` drive
`B200AttnBackend` through these small stand-ins.  They keep the reference's names,
argument meaning and error behaviour for exactly the members the backend reads
(SURVEY.md section 8a rows a3-a5, a10, a11); nothing else of SGLang is rebuilt.
Inside a real SGLang process the backend receives the real objects instead
(see sglang_b200/plugin.py and INTEGRATION.md).
"""
from .attention_registry import ATTENTION_BACKENDS, register_attention_backend
from .base_attn_backend import AttentionBackend
from .forward_batch_info import ForwardBatch, ForwardMode
from .memory_pool import KVWriteLoc, MHATokenToKVPool, ReqToTokenPool
from .model_runner import make_model_runner
from .radix_attention import AttentionType, RadixAttention

__all__ = [
    "ATTENTION_BACKENDS", "register_attention_backend", "AttentionBackend", "ForwardBatch", "ForwardMode",
    "KVWriteLoc", "MHATokenToKVPool", "ReqToTokenPool", "AttentionType", "RadixAttention", "make_model_runner",
]
