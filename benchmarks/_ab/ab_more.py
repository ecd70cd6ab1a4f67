"""A/B of the cases the one-tile / D 256 instances serve, for the library in B200ATTN_LIB."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sweep import extend_case
tag = os.environ.get("B200ATTN_LIB", "default")
for name, a, kw in [
    ("cfg5/16 page 1", (8, 16384, 1024, 32, 8, 128, 5), {}),
    ("4K no prefix", (8, 0, 4096, 32, 8, 128, 5), {}),
    ("verify one-tile", (64, 4096, 8, 32, 8, 128, 5), dict(shared=False, tree_mask=True, one_tile=True)),
    ("D256 gemma", (8, 16384, 1024, 16, 8, 256, 5), {}),
    ("D64 32/8", (8, 16384, 1024, 32, 8, 64, 5), {}),
    ("fp8 e4m3", (8, 16384, 1024, 32, 8, 128, 5), dict(pool_dtype=__import__("torch").float8_e4m3fn)),
]:
    r = extend_case(name, *a, **kw)
    print(tag, name, r["ms"], r["tflops"])
