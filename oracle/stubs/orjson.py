"""Test scaffolding only: json-backed stand-in for `orjson`, which the reference
imports at module load (python/sglang/srt/utils/common.py:81) but this image lacks.
Used solely by oracle/gen_golden.py when importing the reference on CPU."""
import json

OPT_NON_STR_KEYS = OPT_SERIALIZE_NUMPY = OPT_INDENT_2 = OPT_SORT_KEYS = OPT_APPEND_NEWLINE = 0
JSONDecodeError = json.JSONDecodeError


class JSONEncodeError(TypeError):
    pass


def dumps(o, default=None, option=None):
    return json.dumps(o, default=default).encode()


def loads(s):
    return json.loads(s)
