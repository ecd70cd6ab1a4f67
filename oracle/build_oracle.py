"""gcc build of the C restatement of the KV-store byte work (test infrastructure; not part of the product).

    python oracle/build_oracle.py [--force]      ->  oracle/_build/liboracle_kv.so
"""
from __future__ import annotations

import subprocess
import sys
from pathlib import Path

ORACLE_DIR = Path(__file__).resolve().parent
SRC = ORACLE_DIR / "kv_store_ref.c"
OUT = ORACLE_DIR / "_build" / "liboracle_kv.so"


def build_oracle(force: bool = False) -> Path:
    if not force and OUT.exists() and OUT.stat().st_mtime >= SRC.stat().st_mtime:
        return OUT
    OUT.parent.mkdir(parents=True, exist_ok=True)
    subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-o", str(OUT), str(SRC)], check=True)
    return OUT


if __name__ == "__main__":
    print(build_oracle(force="--force" in sys.argv))
