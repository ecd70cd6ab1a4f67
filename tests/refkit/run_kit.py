"""Run the REFERENCE's own attention-backend unit tests against `--attention-backend b200` (child process).

    python tests/refkit/run_kit.py --out report.json [--only substring]

What runs is the reference's unmodified test code staged under `oracle/_ref` (see oracle/build_ref.py):
the test classes the reference ships for its Triton backend
  test/registered/attention/unittests/dense/test_triton.py  TestTritonDenseAttentionBackendCorrectness
  test/registered/attention/unittests/swa/test_triton.py    TestTritonSWAAttentionBackendCorrectness
are subclassed with every `DenseAttentionCase` re-targeted from backend "triton" to "b200"; their test methods
(the dense kit `run_dense_attention_case`, dense_attention.py:1262; the CUDA-graph decode kit
`run_dense_cuda_graph_decode_case`, cuda_graph_decode_runner.py:476; split-op, speculative verify / draft
runners) then drive `B200AttnBackend` through the REAL `ATTENTION_BACKENDS` registry, `ReqToTokenPool`,
`MHATokenToKVPool`, `ForwardBatch`, `RadixAttention` and `ForwardContext`.  The expected values are the kit's own
fp32 reference (dense_attention.py:795-848), tolerance the kit's own 3e-2.

The one thing a maintainer adds on the reference side besides the plugin entry point is the SWA-decode rule
classification of the new backend name (dense_attention.py:53-56): b200 follows the `extend_window` rule
(window + 1 keys, like torch_native / fa3 / trtllm_mha), which this script registers at run time.
"""
from __future__ import annotations

import argparse
import dataclasses
import importlib.util
import json
import os
import sys
import time
import traceback
import unittest
from pathlib import Path

REPO = Path(__file__).resolve().parents[2]
REF = REPO / "oracle" / "_ref"


def _retarget(obj, backend: str):
    """Deep-copy a case container, replacing DenseAttentionCase.backend."""
    if dataclasses.is_dataclass(obj) and hasattr(obj, "backend"):
        return dataclasses.replace(obj, backend=backend)
    if isinstance(obj, tuple):
        return tuple(_retarget(x, backend) for x in obj)
    if isinstance(obj, list):
        return [_retarget(x, backend) for x in obj]
    return obj


def _load(path: Path, name: str):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


class _Result(unittest.TestResult):
    def __init__(self):
        super().__init__()
        self.records = {}

    def _rec(self, test):
        name = f"{type(test).__name__}.{test._testMethodName}"
        return self.records.setdefault(name, {"outcome": "passed", "subtests": 0, "failed_subtests": [],
                                              "detail": "", "seconds": 0.0})

    def startTest(self, test):
        super().startTest(test)
        self._rec(test)["_t0"] = time.time()

    def stopTest(self, test):
        super().stopTest(test)
        r = self._rec(test)
        r["seconds"] = round(time.time() - r.pop("_t0"), 2)

    def addSubTest(self, test, subtest, err):
        super().addSubTest(test, subtest, err)
        r = self._rec(test)
        r["subtests"] += 1
        if err is not None:
            r["outcome"] = "failed"
            r["failed_subtests"].append(str(subtest))
            r["detail"] += f"--- {subtest}\n" + "".join(traceback.format_exception(*err))[-3000:] + "\n"

    def addError(self, test, err):
        super().addError(test, err)
        r = self._rec(test)
        r["outcome"] = "error"
        r["detail"] += "".join(traceback.format_exception(*err))[-4000:]

    def addFailure(self, test, err):
        super().addFailure(test, err)
        r = self._rec(test)
        r["outcome"] = "failed"
        r["detail"] += "".join(traceback.format_exception(*err))[-4000:]

    def addSkip(self, test, reason):
        super().addSkip(test, reason)
        r = self._rec(test)
        r["outcome"] = "skipped"
        r["detail"] = reason


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    report = {"import_error": None, "tests": {}, "sglang_file": None, "base_class": None}

    def finish(rc):
        Path(args.out).write_text(json.dumps(report, indent=1))
        return rc

    if os.environ.get("CUDA_VISIBLE_DEVICES") == "":
        del os.environ["CUDA_VISIBLE_DEVICES"]          # sglang/test/test_utils.py:254 indexes [0]
    sys.path[:0] = [str(REF / "stubs"), str(REF / "python"), str(REPO)]
    try:
        import torch

        import sglang
        from sglang.srt.layers.attention.attention_registry import ATTENTION_BACKENDS
        from sglang.test.kits.attention_unittest.attention_methods import dense_attention as kit

        from sglang_b200 import plugin
        from sglang_b200.backend import B200AttnBackend

        plugin.register()
        assert "b200" in ATTENTION_BACKENDS, "plugin.register() did not reach the real registry"
        kit._SWA_DECODE_EXTEND_WINDOW = frozenset(kit._SWA_DECODE_EXTEND_WINDOW | {"b200"})
        report["sglang_file"] = sglang.__file__
        report["base_class"] = f"{B200AttnBackend.__mro__[1].__module__}.{B200AttnBackend.__mro__[1].__name__}"
        report["cuda"] = torch.cuda.is_available()

        dense = _load(REF / "test" / "attention_unittests" / "dense" / "test_triton.py", "ref_dense_test_triton")
        swa = _load(REF / "test" / "attention_unittests" / "swa" / "test_triton.py", "ref_swa_test_triton")
    except BaseException:
        report["import_error"] = traceback.format_exc()[-6000:]
        return finish(3)

    def derive(base, name):
        ns = {k: _retarget(v, "b200") for k, v in vars(base).items() if k.isupper()}
        ns["CASES"] = _retarget(base.CASES, "b200")
        return type(name, (base,), ns)

    classes = [
        derive(dense.TestTritonDenseAttentionBackendCorrectness, "B200DenseKit"),
        derive(swa.TestTritonSWAAttentionBackendCorrectness, "B200SWAKit"),
    ]
    suite = unittest.TestSuite()
    loader = unittest.TestLoader()
    for cls in classes:
        for name in loader.getTestCaseNames(cls):
            if args.only in f"{cls.__name__}.{name}":
                suite.addTest(cls(name))
    result = _Result()
    suite.run(result)
    report["tests"] = result.records
    bad = [k for k, v in result.records.items() if v["outcome"] in ("failed", "error")]
    print(json.dumps({k: (v["outcome"], v["subtests"], v["failed_subtests"]) for k, v in result.records.items()},
                     indent=1))
    return finish(1 if bad else 0)


if __name__ == "__main__":
    sys.exit(main())
