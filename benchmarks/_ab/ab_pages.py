"""extend cfg5/16 at page sizes 1/16/64 (+ 4K no-prefix prefill) for the library in B200ATTN_LIB."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sweep import extend_case
tag = os.environ.get("B200ATTN_LIB", "default")
for page in (1, 16, 64):
    r = extend_case("cfg5/16", 8, 16384, 1024, 32, 8, 128, 5, page=page)
    print(tag, "page", page, r["ms"], r["tflops"])
r = extend_case("4K", 8, 0, 4096, 32, 8, 128, 5)
print(tag, "prefill 4K no prefix", r["ms"], r["tflops"])
