"""Per-warp P hand-off times of one extend_tc_kernel CTA (-DB200_DEV -DB200_TRACE=<block> -DB200_TRACE_WARPS build)."""
import ctypes, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from sweep import extend_case
from sglang_b200 import _C
print(extend_case("trace 8 req x (16K + 1K)", 8, 16384, 1024, 32, 8, 128, 3))
torch.cuda.synchronize()
buf = np.zeros(16 * 512, dtype=np.int64)
assert _C.lib.b200attn_dev_trace_read(buf.ctypes.data_as(ctypes.c_void_p)) == 0
tr = buf.reshape(16, 512)
lo, hi = 20, 120
NW = int((tr[:, lo] > 0).sum()) // 2      # softmax warps per tile (4: row per thread, 8: two threads per row)
for t in (0, 1):
    w = tr[NW * t:NW * t + NW, lo:hi]
    first = w.min(axis=0)
    print(f"tile {t}: mean lag of each warp's P hand-off behind the tile's first warp (warp = 8t + 4h + quarter):")
    print("   ", np.round((w - first).mean(axis=1), 0))
    print("    last-arriver histogram:", np.bincount(w.argmax(axis=0), minlength=NW))
