"""The API server's JSON reader (csrc/host/json.hpp) against Python's json module: property test over random documents
(hypothesis), plus the rejections that matter for a network-facing parser."""
import ctypes as C
import json
import math

import pytest

pytest.importorskip("hypothesis")
from hypothesis import given, settings, strategies as st  # noqa: E402

from dllama_b200.text.native import host_lib


def _roundtrip(raw: bytes):
    L = host_lib()
    L.dl_json_roundtrip.restype = C.c_int
    L.dl_json_roundtrip.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int]
    cap = 6 * len(raw) + 64
    out = C.create_string_buffer(cap)
    n = L.dl_json_roundtrip(raw, len(raw), out, cap)
    return None if n < 0 else out.raw[:n]


_leaf = st.one_of(st.none(), st.booleans(), st.integers(-2 ** 62, 2 ** 62), st.floats(allow_nan=False, allow_infinity=False),
                  st.text(max_size=20))
_doc = st.recursive(_leaf, lambda kids: st.one_of(st.lists(kids, max_size=5), st.dictionaries(st.text(max_size=8), kids, max_size=5)),
                    max_leaves=25)


@settings(max_examples=300, deadline=None)
@given(_doc, st.booleans())
def test_reader_agrees_with_python_json(doc, ascii_only):
    raw = json.dumps(doc, ensure_ascii=ascii_only).encode("utf-8")        # \\uXXXX escapes (surrogate pairs) or raw UTF-8
    back = _roundtrip(raw)
    assert back is not None, raw
    got = json.loads(back.decode("utf-8"))

    def same(a, b):
        if isinstance(a, float) or isinstance(b, float):
            return isinstance(b, (int, float)) and (a == b or math.isclose(a, b, rel_tol=1e-15))
        if isinstance(a, dict):
            return isinstance(b, dict) and a.keys() == b.keys() and all(same(a[k], b[k]) for k in a)
        if isinstance(a, list):
            return isinstance(b, list) and len(a) == len(b) and all(same(x, y) for x, y in zip(a, b))
        return a == b and type(a) is type(b)
    assert same(doc, got), (doc, got)


@pytest.mark.parametrize("raw", [b"", b"{", b"[1,]", b"{\"a\":}", b"{\"a\" 1}", b"[1 2]", b"\"abc", b"\"\\x\"", b"\"\\u12\"", b"tru",
                                 b"nul", b"01x", b"-", b"1 2", b"{\"a\":1}}", b"\"a\nb\"", b"[" * 100 + b"]" * 100, b"\xff"])
def test_reader_rejects_malformed_documents(raw):
    if raw == b"\xff":                       # bytes outside a string are a syntax error; inside a string they are kept and
        assert _roundtrip(raw) is None       # replaced by U+FFFD when the text is written back out
        assert _roundtrip(b"\"\xff\"") == "\"\ufffd\"".encode("utf-8")
    else:
        assert _roundtrip(raw) is None, raw


def test_reader_edge_cases():
    assert json.loads(_roundtrip(b" [ 1e3 , -0.5 , 12345678901234567890 , \"\\ud83d\\ude00\" , \"\\ud800\" ] ")) == \
        [1000.0, -0.5, 12345678901234567890.0, "\U0001F600", "\ufffd"]
    assert _roundtrip(b"[" * 60 + b"]" * 60) is not None            # nesting up to 64 levels is accepted
