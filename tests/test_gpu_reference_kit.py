"""Acceptance test: the reference's OWN backend unit tests, driven against `--attention-backend b200`.

`tests/refkit/run_kit.py` runs once per session in a child process (the staged reference `oracle/_ref`
on its path, `sglang_b200.plugin.register()` into the real `ATTENTION_BACKENDS`) and reports per test
method; each method of the reference's Triton test classes is one pytest case here.
Reference: test/registered/attention/unittests/dense/test_triton.py, swa/test_triton.py,
sglang/test/kits/attention_unittest/attention_methods/dense_attention.py:996,1262,
runner_modes/cuda_graph_decode_runner.py:289-322,476.
"""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
REF = REPO / "oracle" / "_ref"

pytestmark = pytest.mark.gpu

DENSE = [
    "test_projected_dense_attention_cases",
    "test_layout_robustness_cases",
    "test_runner_mode_cuda_graph_decode_cases",
    "test_runner_mode_split_op_extend_cases",
    "test_runner_mode_spec_verify_cases",
    "test_runner_mode_spec_verify_cuda_graph_cases",
    "test_runner_mode_eagle_draft_extend_v2_cuda_graph_cases",
    "test_runner_mode_eagle_draft_extend_v2_cuda_graph_runner_cases",
    "test_runner_mode_eagle_draft_cuda_graph_runner_cases",
    "test_runner_mode_frozen_kv_mtp_cuda_graph_runner_cases",
]
SWA = [
    "test_projected_swa_attention_cases",
    "test_layout_robustness_cases",
    "test_runner_mode_cuda_graph_decode_cases",
    "test_runner_mode_split_op_extend_cases",
    "test_runner_mode_spec_verify_cases",
    "test_runner_mode_spec_verify_cuda_graph_cases",
]
# The multi-step EAGLE *draft* runners build their per-step draft backends through DraftBackendFactory, whose
# backend-name -> class maps are dict literals inside the methods (srt/speculative/draft_utils.py:71-88, 100-118).
# sglang_b200.plugin.register() patches DraftBackendFactory._create_backend for the name "b200": draft-extend ->
# B200AttnBackend, chain draft decode (topk 1) -> B200MultiStepDraftBackend, tree draft decode -> the reference's own
# Triton wrapper.  All 16 methods run.
ALL = [f"B200DenseKit.{n}" for n in DENSE] + [f"B200SWAKit.{n}" for n in SWA]


@pytest.fixture(scope="session")
def kit_report(tmp_path_factory):
    if not (REF / "MANIFEST.json").exists():
        pytest.skip("oracle/_ref is not staged (run `python oracle/build_ref.py` in the build container)")
    out = tmp_path_factory.mktemp("refkit") / "report.json"
    env = dict(os.environ, PYTHONDONTWRITEBYTECODE="1")
    proc = subprocess.run([sys.executable, str(REPO / "tests" / "refkit" / "run_kit.py"), "--out", str(out)],
                          env=env, capture_output=True, text=True, timeout=1500)
    if not out.exists():
        pytest.fail(f"reference kit runner died (rc {proc.returncode}):\n{proc.stderr[-4000:]}")
    report = json.loads(out.read_text())
    keep = REPO / "gpurun_out"
    if keep.is_dir():
        (keep / "refkit_report.json").write_text(json.dumps(report, indent=1))
    return report


def test_reference_tree_imports_and_sees_plugin(kit_report):
    assert kit_report["import_error"] is None, kit_report["import_error"]
    assert "oracle/_ref/python/sglang" in kit_report["sglang_file"]
    # B200AttnBackend subclasses the reference's real ABC inside an SGLang process
    assert kit_report["base_class"] == "sglang.srt.layers.attention.base_attn_backend.AttentionBackend"
    assert kit_report["cuda"] is True


@pytest.mark.parametrize("name", ALL)
def test_reference_kit(kit_report, name):
    assert kit_report["import_error"] is None, kit_report["import_error"]
    rec = kit_report["tests"].get(name)
    assert rec is not None, f"{name} did not run; ran: {sorted(kit_report['tests'])}"
    if rec["outcome"] == "skipped":
        pytest.skip(rec["detail"])
    # (sglang_b200.plugin teaches DraftBackendFactory the plugin's name, so the draft-runner cases run too)
    assert "EAGLE is not supported in" not in rec["detail"], rec["detail"][-2000:]
    assert rec["outcome"] == "passed", f"{rec['failed_subtests']}\n{rec['detail'][-6000:]}"
