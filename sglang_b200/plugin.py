"""SGLang plugin entry point: registers `--attention-backend b200`.

Reference hooks used (python/sglang/srt/...):
  layers/attention/attention_registry.py:31-39   register_attention_backend(name)
  server_args.py:386-387                         add_attention_backend_choices([...])
  plugins/__init__.py:29,103-141                 entry-point group `sglang.srt.plugins`
`load_plugins()` runs before argument parsing and again in every scheduler process
(launch_server.py:64-68, managers/scheduler.py:4894-4895), so this function is idempotent.
"""
from __future__ import annotations

from . import BACKEND_NAME


def create_b200_backend(runner):
    from .backend import B200AttnBackend

    return B200AttnBackend(runner)


def register() -> None:
    """Install the backend into a running SGLang process (or into the mirror registry)."""
    try:
        from sglang.srt.layers.attention.attention_registry import (  # type: ignore
            ATTENTION_BACKENDS,
            register_attention_backend,
        )
    except Exception:
        from .srt_mirror.attention_registry import ATTENTION_BACKENDS, register_attention_backend

    if BACKEND_NAME not in ATTENTION_BACKENDS:
        register_attention_backend(BACKEND_NAME)(create_b200_backend)
    try:
        from sglang.srt.server_args import add_attention_backend_choices  # type: ignore

        add_attention_backend_choices([BACKEND_NAME])
    except Exception:
        pass
