"""Full cfg5 (128 requests x (16 K shared prefix + 1 K new), page 1 and 64) for the library in B200ATTN_LIB: the
power-capped case, where MUFU / FMA-pipe trade-offs differ from the 8-request case the kernel was tuned on."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sweep import extend_case
tag = os.environ.get("B200ATTN_LIB", "default")
for page in (1, 64):
    r = extend_case("cfg5", 128, 16384, 1024, 32, 8, 128, 3, page=page)
    print(tag, "full cfg5 page", page, r["ms"], r["tflops"], flush=True)
