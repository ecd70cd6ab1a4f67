"""One speculative-verify launch pair (split-KV attention + merge) for profiling:
python benchmarks/verify_split_one.py [bs ctx draft splits(0=auto)]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent))
from sweep import extend_case
a = [int(x) for x in sys.argv[1:]] + [None] * 4
bs, ctx, n, sp = a[0] or 1, a[1] or 32768, a[2] or 8, a[3] if a[3] is not None else 0
print(extend_case(f"verify {bs} req x ({ctx} + {n} draft, tree mask)", bs, ctx, n, 32, 8, 128, 5, shared=False, tree_mask=True,
                  one_tile=True, splits=sp))
