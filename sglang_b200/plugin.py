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
    _patch_draft_backend_factory()


def _patch_draft_backend_factory() -> None:
    """Speculative decoding: `DraftBackendFactory` resolves the draft model's backends through dicts that are local to
    its methods (speculative/draft_utils.py:60-135), so a plugin backend name ends in "EAGLE is not supported in
    attention backend b200".  Teach `_create_backend` the name: the draft-extend backend is a `B200AttnBackend`, the
    multi-step draft-decode backend is `B200MultiStepDraftBackend` (chain and tree drafting)."""
    try:
        from sglang.srt.speculative.draft_utils import DraftBackendFactory  # type: ignore
    except Exception:
        return
    if getattr(DraftBackendFactory, "_b200_patched", False):
        return
    original = DraftBackendFactory._create_backend

    def _create_backend(self, backend_name, backend_map, error_template):
        backend_type = self.draft_attn_backend or getattr(self.server_args, backend_name, None)
        if backend_type is None:
            backend_type = self.server_args.attention_backend
        if backend_type != BACKEND_NAME:
            return original(self, backend_name, backend_map, error_template)
        from .backend import B200AttnBackend, B200MultiStepDraftBackend

        is_decode_map = getattr(backend_map.get("triton"), "__name__", "").endswith("_decode_backend")
        if not is_decode_map:
            return B200AttnBackend(self.draft_model_runner, skip_prefill=False)
        return B200MultiStepDraftBackend(self.draft_model_runner, self.topk, self.speculative_num_steps)

    DraftBackendFactory._create_backend = _create_backend
    DraftBackendFactory._b200_patched = True
