"""Child process: import the REAL reference classes (oracle/_ref) beside the mirrors of sglang_b200/srt_mirror and
report, member by member, whether what `B200AttnBackend` reads exists in the reference with the same meaning.
Runs on CPU (no compute).  Output: one JSON object on the last stdout line."""
from __future__ import annotations

import dataclasses
import inspect
import json
import os
import sys
from pathlib import Path
from types import SimpleNamespace

REPO = Path(__file__).resolve().parents[2]
REF = REPO / "oracle" / "_ref"
if os.environ.get("CUDA_VISIBLE_DEVICES") == "":
    del os.environ["CUDA_VISIBLE_DEVICES"]
sys.path[:0] = [str(REF / "stubs"), str(REF / "python"), str(REPO)]

import torch  # noqa: E402

problems = []
checked = []


def check(cond, what):
    checked.append(what)
    if not cond:
        problems.append(what)


from sglang.srt.layers.attention import attention_registry as real_reg  # noqa: E402
from sglang.srt.layers.attention.base_attn_backend import AttentionBackend as RealABC  # noqa: E402
from sglang.srt.layers.radix_attention import AttentionType as RealAT  # noqa: E402
from sglang.srt.layers.radix_attention import RadixAttention as RealRA  # noqa: E402
from sglang.srt.mem_cache.memory_pool import KVWriteLoc as RealLoc  # noqa: E402
from sglang.srt.mem_cache.memory_pool import MHATokenToKVPool as RealPool  # noqa: E402
from sglang.srt.mem_cache.memory_pool import ReqToTokenPool as RealRTP  # noqa: E402
from sglang.srt.model_executor.forward_batch_info import ForwardBatch as RealFB  # noqa: E402
from sglang.srt.model_executor.forward_batch_info import ForwardMode as RealFM  # noqa: E402
from sglang.srt.plugins import GENERAL_PLUGINS_GROUP  # noqa: E402

from sglang_b200 import plugin, srt_mirror as M  # noqa: E402
from sglang_b200.backend import B200AttnBackend  # noqa: E402

# ---- ABC: the hooks the mirror declares exist in the reference with the same parameter names
for name in ("init_forward_metadata", "init_forward_metadata_out_graph", "init_forward_metadata_in_graph",
             "init_cuda_graph_state", "get_cuda_graph_seq_len_fill_value", "on_after_cuda_graph_warmup",
             "forward", "forward_decode", "forward_extend", "support_triton",
             "update_verify_buffers_to_fill_after_draft", "draft_extend_metadata_captured_in_graph"):
    real = getattr(RealABC, name, None)
    check(real is not None, f"ABC.{name} exists in the reference")
    if real is None:
        continue
    rp = [p for p in inspect.signature(real).parameters if p not in ("self", "kwargs")]
    mp = [p for p in inspect.signature(getattr(M.AttentionBackend, name)).parameters if p not in ("self", "kwargs")]
    bp = [p for p in inspect.signature(getattr(B200AttnBackend, name)).parameters if p not in ("self", "kwargs")]
    check(rp == mp, f"ABC.{name} parameters: reference {rp} == mirror {mp}")
    check(bp[: len(rp)] == rp, f"B200AttnBackend.{name} accepts the reference's parameters {rp} (has {bp})")
check(issubclass(B200AttnBackend, RealABC), "B200AttnBackend subclasses the reference ABC inside an SGLang process")
# hooks whose ABC default raises must be overridden when a stock runner / speculative worker calls them on every
# backend (decode_cuda_graph_runner.py:318, eagle_worker_common.py:527)
for name in ("init_cuda_graph_state", "get_cuda_graph_seq_len_fill_value", "forward_decode", "forward_extend",
             "update_verify_buffers_to_fill_after_draft"):
    check(getattr(B200AttnBackend, name) is not getattr(RealABC, name), f"B200AttnBackend overrides ABC.{name}")
for name in ("supports_ragged_verify_graph", "use_captured_forward_metadata_for_breakable_cuda_graph",
             "supports_full_cuda_graph_chunked_prefix", "needs_cpu_seq_lens", "verify_mask"):
    check(hasattr(RealABC, name) and hasattr(M.AttentionBackend, name), f"ABC.{name} exists in the reference and the mirror")

# ---- ForwardMode: same member names and the predicates agree on every member
for m in M.ForwardMode:
    check(hasattr(RealFM, m.name), f"ForwardMode.{m.name} exists in the reference")
preds = ("is_extend", "is_prefill", "is_decode", "is_mixed", "is_idle", "is_decode_or_idle", "is_target_verify",
         "is_draft_extend_v2", "is_cuda_graph")
for m in M.ForwardMode:
    if not hasattr(RealFM, m.name):
        continue
    r = getattr(RealFM, m.name)
    for pr in preds:
        check(getattr(m, pr)() == getattr(r, pr)(), f"ForwardMode.{m.name}.{pr}() agrees")
    check(m.is_extend(include_draft_extend_v2=True) == r.is_extend(include_draft_extend_v2=True),
          f"ForwardMode.{m.name}.is_extend(include_draft_extend_v2=True) agrees")
check({m.name for m in RealFM} == {m.name for m in M.ForwardMode}, "ForwardMode member sets are equal")

# ---- ForwardBatch: every mirror field is a reference field
real_fields = {f.name for f in dataclasses.fields(RealFB)}
for f in dataclasses.fields(M.ForwardBatch):
    # (mixin members such as mha_return_lse, forward_batch_deepseek_mha_mixin.py:50, are class attributes)
    check(f.name in real_fields or hasattr(RealFB, f.name), f"ForwardBatch.{f.name} exists in the reference")

# ---- pools: construct both on CPU and compare what the backend reads
kw = dict(size=64, page_size=16, head_num=2, head_dim=8, layer_num=2, device="cpu", enable_memory_saver=False)
for dt in (torch.bfloat16, torch.float8_e4m3fn):
    rp_ = RealPool(dtype=dt, enable_alt_stream=False, **kw)
    mp_ = M.MHATokenToKVPool(dtype=dt, **kw)
    for attr in ("size", "page_size", "dtype", "store_dtype", "head_num", "head_dim", "layer_num", "start_layer"):
        check(getattr(rp_, attr) == getattr(mp_, attr), f"MHATokenToKVPool[{dt}].{attr}: "
              f"{getattr(rp_, attr)} == {getattr(mp_, attr)}")
    check(isinstance(rp_.k_buffer, list) and len(rp_.k_buffer) == len(mp_.k_buffer), "k_buffer is a per-layer list")
    for buf in ("k_buffer", "v_buffer"):
        a, b = getattr(rp_, buf)[0], getattr(mp_, buf)[0]
        check(a.shape == b.shape and a.dtype == b.dtype and a.stride() == b.stride(),
              f"MHATokenToKVPool[{dt}].{buf}[0] {tuple(a.shape)} {a.dtype} == mirror {tuple(b.shape)} {b.dtype}")
    for fn in ("get_key_buffer", "get_value_buffer"):
        a, b = getattr(rp_, fn)(1), getattr(mp_, fn)(1)
        check(a.shape == b.shape and a.dtype == b.dtype, f"MHATokenToKVPool[{dt}].{fn}(1) view {a.dtype} == {b.dtype}")
    sp = list(inspect.signature(RealPool.set_kv_buffer).parameters)
    mpn = list(inspect.signature(M.MHATokenToKVPool.set_kv_buffer).parameters)
    check(sp[:6] == mpn[:6], f"set_kv_buffer parameters {sp[:6]} == mirror {mpn[:6]}")
check([f.name for f in dataclasses.fields(RealLoc)][:2] == ["loc", "swa_loc"], "KVWriteLoc(loc, swa_loc, ...)")
rr = RealRTP(size=5, max_context_len=33, device="cpu", enable_memory_saver=False)
mr = M.ReqToTokenPool(size=5, max_context_len=33, device="cpu")
check(rr.req_to_token.shape == mr.req_to_token.shape and rr.req_to_token.dtype == mr.req_to_token.dtype,
      f"ReqToTokenPool.req_to_token {tuple(rr.req_to_token.shape)} {rr.req_to_token.dtype} == mirror "
      f"{tuple(mr.req_to_token.shape)} {mr.req_to_token.dtype}")
check(rr.size == mr.size and rr.max_context_len == mr.max_context_len, "ReqToTokenPool.size / max_context_len")

# ---- RadixAttention: the per-layer fields
ra = RealRA(num_heads=8, head_dim=16, scaling=0.25, num_kv_heads=2, layer_id=3, logit_cap=30.0,
            sliding_window_size=7)
ma = M.RadixAttention(num_heads=8, head_dim=16, scaling=0.25, num_kv_heads=2, layer_id=3, logit_cap=30.0,
                      sliding_window_size=7)
for attr in ("tp_q_head_num", "tp_k_head_num", "tp_v_head_num", "qk_head_dim", "v_head_dim", "scaling", "layer_id",
             "logit_cap", "sliding_window_size", "is_cross_attention", "k_scale", "v_scale", "k_scale_float",
             "v_scale_float"):
    check(hasattr(ra, attr), f"RadixAttention.{attr} exists in the reference")
    if hasattr(ra, attr):
        check(getattr(ra, attr) == getattr(ma, attr), f"RadixAttention.{attr}: {getattr(ra, attr)!r} == "
              f"{getattr(ma, attr)!r}")
check(ra.attn_type.value == ma.attn_type.value, "RadixAttention.attn_type default value")
check({a.value for a in M.AttentionType} <= {a.value for a in RealAT}, "AttentionType values exist in the reference")
ra0 = RealRA(num_heads=8, head_dim=16, scaling=0.25, num_kv_heads=2, layer_id=0)
ma0 = M.RadixAttention(num_heads=8, head_dim=16, scaling=0.25, num_kv_heads=2, layer_id=0)
check(ra0.sliding_window_size == ma0.sliding_window_size, "RadixAttention.sliding_window_size default "
      f"{ra0.sliding_window_size} == {ma0.sliding_window_size}")

# ---- registry / plugin
plugin.register()
check("b200" in real_reg.ATTENTION_BACKENDS, "plugin.register() reaches the real ATTENTION_BACKENDS")
from sglang.srt import server_args as real_sa  # noqa: E402

check("b200" in real_sa.ATTENTION_BACKEND_CHOICES, "plugin.register() extends ATTENTION_BACKEND_CHOICES")
check(GENERAL_PLUGINS_GROUP == "sglang.srt.plugins", "entry-point group name")
try:
    import tomllib

    eps = tomllib.loads((REPO / "pyproject.toml").read_text())["project"]["entry-points"]
    check(eps.get(GENERAL_PLUGINS_GROUP, {}).get("b200_attention") == "sglang_b200.plugin:register",
          "pyproject.toml declares the entry point in the reference's group")
except Exception as exc:  # pragma: no cover
    check(False, f"pyproject.toml entry point readable ({exc})")
# construction on a CPU runner must fail loudly (no CPU fallback), with an error type the kit treats as fatal
try:
    real_reg.ATTENTION_BACKENDS["b200"](SimpleNamespace(device="cpu"))
    check(False, "constructing the backend on a CPU device raises")
except RuntimeError:
    check(True, "constructing the backend on a CPU device raises RuntimeError")

print(json.dumps({"checked": len(checked), "problems": problems}))
