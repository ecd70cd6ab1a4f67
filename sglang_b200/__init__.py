"""sglang_b200: a B200-native (sm_100a) attention backend for SGLang's RadixAttention hot path.

    KV-store scatter  ->  paged split-KV decode  /  prefix + causal extend

Hand-written CUDA behind a C ABI (include/b200attn.h, libb200attn.so), driven through the
reference's AttentionBackend plugin API (`sglang_b200.backend.B200AttnBackend`).
`sglang_b200.plugin.register` installs it into an SGLang process as `--attention-backend b200`.
"""
__version__ = "0.1.0"

BACKEND_NAME = "b200"

__all__ = ["BACKEND_NAME", "__version__"]
