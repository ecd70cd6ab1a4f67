"""CPU contract test: the reference's REAL classes (staged under oracle/_ref by oracle/build_ref.py) against the
stand-ins in sglang_b200/srt_mirror and the members B200AttnBackend reads (SURVEY section 8a rows a1-a5, a10, a11,
section 8b).  The GPU counterpart is tests/test_gpu_reference_kit.py (the reference's own kit driving the backend)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent


def test_mirror_matches_reference_classes():
    if not (REPO / "oracle" / "_ref" / "MANIFEST.json").exists():
        pytest.skip("oracle/_ref is not staged (run `python oracle/build_ref.py` in the build container)")
    proc = subprocess.run([sys.executable, str(REPO / "tests" / "refkit" / "contract_probe.py")],
                          env=dict(os.environ, PYTHONDONTWRITEBYTECODE="1"), capture_output=True, text=True,
                          timeout=600)
    assert proc.returncode == 0, proc.stderr[-4000:]
    report = json.loads(proc.stdout.strip().splitlines()[-1])
    assert report["checked"] > 150
    assert report["problems"] == [], "\n".join(report["problems"])
