"""One TARGET_VERIFY-shaped extend (tree mask) for profiling: python benchmarks/verify_one.py [bs P E]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent))
from sweep import extend_case
a = [int(x) for x in sys.argv[1:]] + [None] * 3
bs, P, E = a[0] or 64, a[1] or 4096, a[2] or 8
print(extend_case(f"verify {bs} req x ({P} + {E} draft, tree mask)", bs, P, E, 32, 8, 128, 3, shared=False, tree_mask=True))
