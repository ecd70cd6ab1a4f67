"""CPU contract of sglang_b200/adapters.py: same field / parameter names as the reference's own definitions, read
from the reference sources with `ast` (no import).  Skipped where /root/reference does not exist (the GPU box)."""
import ast
import inspect
from pathlib import Path

import pytest

REF = Path("/root/reference/python/sglang")
pytestmark = pytest.mark.skipif(not REF.exists(), reason="reference checkout not present")


def _find(tree, kind, name):
    for node in ast.walk(tree):
        if isinstance(node, kind) and node.name == name:
            return node
    raise AssertionError(f"{name} not found")


def _params(fn):
    a = fn.args
    return [x.arg for x in a.posonlyargs + a.args], [x.arg for x in a.kwonlyargs]


def test_fused_set_kv_buffer_arg_and_rope_signature():
    from dataclasses import fields

    from sglang_b200 import adapters

    tree = ast.parse((REF / "kernels/ops/attention/rope.py").read_text())
    cls = _find(tree, ast.ClassDef, "FusedSetKVBufferArg")
    ref_fields = [n.target.id for n in cls.body if isinstance(n, ast.AnnAssign)]
    assert [f.name for f in fields(adapters.FusedSetKVBufferArg)] == ref_fields
    pos, kw = _params(_find(tree, ast.FunctionDef, "apply_rope_with_cos_sin_cache_inplace"))
    sig = inspect.signature(adapters.apply_rope_with_cos_sin_cache_inplace)
    ours_pos = [n for n, p in sig.parameters.items() if p.kind == p.POSITIONAL_OR_KEYWORD]
    ours_kw = [n for n, p in sig.parameters.items() if p.kind == p.KEYWORD_ONLY]
    assert ours_pos == pos and ours_kw == kw


def test_create_fused_arg_and_row_parallel_forward_signature():
    from sglang_b200 import adapters

    tree = ast.parse((REF / "srt/models/utils.py").read_text())
    pos, _ = _params(_find(tree, ast.FunctionDef, "create_fused_set_kv_buffer_arg"))
    assert list(inspect.signature(adapters.create_fused_set_kv_buffer_arg).parameters) == pos

    tree = ast.parse((REF / "srt/layers/linear.py").read_text())
    cls = _find(tree, ast.ClassDef, "RowParallelLinear")
    fwd = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "forward")
    ref_pos, _ = _params(fwd)
    ours = list(inspect.signature(adapters.B200RowParallelLinear.forward).parameters)
    # every parameter of ours exists in the reference with the same name, in the reference's order
    assert ours[0] == ref_pos[0] == "self" and ours[1] == ref_pos[1] == "input_"
    assert [p for p in ref_pos if p in ours] == ours
    init = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "__init__")
    ref_init = set(_params(init)[0]) | set(_params(init)[1])
    for name in ("input_size", "output_size", "bias", "input_is_parallel", "skip_bias_add", "params_dtype",
                 "reduce_results", "tp_rank", "tp_size"):
        assert name in ref_init and name in inspect.signature(adapters.B200RowParallelLinear.__init__).parameters
