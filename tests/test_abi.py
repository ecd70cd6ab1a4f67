"""The C-ABI library loads and exports every symbol include/b200attn.h declares (no GPU needed)."""
import ctypes
import re
from pathlib import Path

import pytest
import torch

REPO = Path(__file__).resolve().parent.parent


def _declared_symbols():
    text = (REPO / "include" / "b200attn.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200attn_\w+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    from sglang_b200 import _C

    declared = _declared_symbols()
    assert len(declared) >= 12
    for name in declared:
        assert hasattr(_C.lib, name), f"{name} declared in b200attn.h but not exported"
    assert sorted(_C.SYMBOLS) == declared, "ctypes table and header disagree"
    assert _C.lib.b200attn_abi_version() == _C.ABI_VERSION


def test_struct_layouts_match_header_field_order():
    """Field names of the ctypes structures appear in the header in the same order."""
    from sglang_b200 import _C

    text = (REPO / "include" / "b200attn.h").read_text()
    for cname, cls in [("b200_store_kv_args", _C.StoreKVArgs), ("b200_decode_plan_args", _C.DecodePlanArgs),
                       ("b200_decode_args", _C.DecodeArgs), ("b200_extend_plan_args", _C.ExtendPlanArgs),
                       ("b200_extend_args", _C.ExtendArgs)]:
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (cname, cname), text, flags=re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            for part in decl.split(","):
                names.append(re.findall(r"(\w+)\s*$", part.strip())[0])
        assert names == [f[0] for f in cls._fields_], cname


def test_pure_host_queries_work_without_gpu():
    from sglang_b200 import _C, ops

    assert _C.lib.b200attn_decode_plan_ints(16, 100) == 8 + 17 + 100
    assert _C.lib.b200attn_extend_plan_ints(16, 100) == 8 + 17 + 200
    assert ops.decode_fast_supported(128, 128, 4, torch.bfloat16, torch.bfloat16)
    assert ops.decode_fast_supported(128, 128, 8, torch.bfloat16, torch.float8_e4m3fn)
    assert not ops.decode_fast_supported(80, 80, 1, torch.bfloat16, torch.bfloat16)
    assert not ops.decode_fast_supported(128, 128, 32, torch.bfloat16, torch.bfloat16)
    # tcgen05 extend instances: D 128 and D 64, bf16 / f16 / FP8 (e4m3, e5m2) pools; everything else is the generic kernel
    for D in (64, 128):
        assert ops.extend_fast_supported(D, D, 4, torch.bfloat16, torch.bfloat16)
        assert ops.extend_fast_supported(D, D, 7, torch.float16, torch.float8_e4m3fn)
        assert ops.extend_block_q(D, D, 4, torch.bfloat16, torch.bfloat16) == 64   # 2 tiles x 128 / group tokens
    assert ops.extend_fast_supported(256, 256, 4, torch.bfloat16, torch.bfloat16)        # one-tile D 256 instance
    assert ops.extend_fast_supported(256, 256, 4, torch.bfloat16, torch.float8_e4m3fn)   # ... also with FP8 pools
    assert ops.extend_fast_supported(256, 256, 4, torch.float16, torch.float8_e5m2)
    assert ops.extend_block_q(256, 256, 4, torch.bfloat16, torch.bfloat16) == 32         # one 128-row tile
    assert ops.extend_block_q_one_tile(128, 128, 4, torch.bfloat16, torch.bfloat16) == 32
    assert ops.decode_fast_supported(128, 128, 4, torch.bfloat16, torch.float8_e5m2)
    assert not ops.extend_fast_supported(128, 64, 4, torch.bfloat16, torch.bfloat16)
    assert ops.extend_fast_supported(128, 128, 4, torch.bfloat16, torch.float8_e5m2)


def test_ops_refuse_cpu_tensors():
    """No CPU fallback: the product path fails loudly instead of computing on the host."""
    from sglang_b200 import ops

    k = torch.zeros(2, 8, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.store_kv(k, k, torch.zeros(4, 8, dtype=torch.bfloat16), torch.zeros(4, 8, dtype=torch.bfloat16),
                     torch.tensor([1, 2]))


def test_bad_arguments_return_error_codes():
    from sglang_b200 import _C

    assert _C.lib.b200attn_store_kv(None, None) == -1
    assert b"null" in _C.lib.b200attn_last_error()
    a = _C.DecodePlanArgs(seq_lens=1, plan=1, batch_size=4, max_bs=2, max_chunks=10, num_kv_heads=1,
                          num_ctas=1, waves=1, min_chunk_tokens=64)
    assert _C.lib.b200attn_decode_plan(ctypes.byref(a), None) == -1
    assert b"max_bs" in _C.lib.b200attn_last_error()


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    import importlib
    import sys

    monkeypatch.setenv("B200ATTN_LIB", str(tmp_path / "nope.so"))
    saved = sys.modules.pop("sglang_b200._C")
    try:
        with pytest.raises(ImportError, match="no CPU or PyTorch fallback"):
            importlib.import_module("sglang_b200._C")
    finally:
        sys.modules["sglang_b200._C"] = saved
