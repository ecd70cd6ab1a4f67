"""Name -> factory registry (reference srt/layers/attention/attention_registry.py:31-39)."""
ATTENTION_BACKENDS = {}


def register_attention_backend(name):
    def decorator(fn):
        ATTENTION_BACKENDS[name] = fn
        return fn

    return decorator
