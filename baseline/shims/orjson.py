"""stdlib-json stand-in for ``orjson`` (reference: data_utils/fed_emnist.py:1)."""
import json as _json
import _ref_compat  # noqa: F401


def loads(s):
    return _json.loads(s)


def dumps(obj):
    return _json.dumps(obj).encode("utf-8")
