"""In-tree nvcc build of libb200attn.so (sm_100a only).

    python -m sglang_b200.build            # build if sources are newer than the .so
    python -m sglang_b200.build --force

The .so lands in sglang_b200/lib/ (git-ignored, shipped to the GPU box by gpurun).
"""
from __future__ import annotations

import argparse
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO_DIR = PKG_DIR.parent
CSRC = PKG_DIR / "csrc"
LIB_DIR = PKG_DIR / "lib"
LIB_PATH = LIB_DIR / "libb200attn.so"

CU_SOURCES = ["abi.cu", "kv_store.cu", "decode.cu", "extend.cu", "merge_state.cu", "oproj_allreduce.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libb200attn cannot be built")


def _stale(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    deps = sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.inc")) + [REPO_DIR / "include" / "b200attn.h"]
    if not force and not _stale(LIB_PATH, deps):
        return LIB_PATH
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    objs = []
    procs = []
    nvcc = _nvcc()
    for src in CU_SOURCES:
        obj = LIB_DIR / (src + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(REPO_DIR / "include"), "-c", str(CSRC / src), "-o", str(obj)]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    log = []
    for src, proc in procs:
        out, _ = proc.communicate()
        log.append(f"==== {src}\n{out}")
        if proc.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{out}")
    (LIB_DIR / "ptxas.log").write_text("\n".join(log))
    cmd = [nvcc, "-shared", "-Wno-deprecated-gpu-targets", "-o", str(LIB_PATH), *map(str, objs)]
    subprocess.run(cmd, check=True)
    if verbose:
        print(f"built {LIB_PATH}")
    return LIB_PATH


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    args = ap.parse_args(argv)
    build_cuda(force=args.force, verbose=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
