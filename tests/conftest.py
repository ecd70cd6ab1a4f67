import os
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
if str(REPO) not in sys.path:
    sys.path.insert(0, str(REPO))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """Build libb200attn.so and the C oracle once per session (nvcc cross-compiles without a GPU)."""
    from oracle.build_oracle import build_oracle
    from sglang_b200.build import build_cuda

    build_cuda()
    build_oracle()
