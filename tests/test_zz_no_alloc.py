"""dl_argmax_noalloc through the C ABI.

Some users of this library cannot allocate device memory at all: they run as a plugin inside an
application that owns the memory budget, or in a serving loop where allocator calls are banned once
traffic starts, or they simply have raw pointers and no allocator.  For them the library offers a
greedy-pick entry point whose scratch space is a fixed array inside the library.  This test calls it
the way such a caller would -- plain ctypes, raw device pointers, no torch op in between -- and
checks, twice in a row, that it picks exactly the tokens the regular entry point picks.

Collected last (file name) so that a failure here cannot hide the results of the other GPU tests.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from dllama_b200.ops.lib import cuda_lib
    L = cuda_lib()
    L.dl_argmax_noalloc.restype = C.c_int
    L.dl_argmax_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return L


def test_noalloc_argmax_matches_regular_entry_point_twice():
    from dllama_b200.ops import kernels as K
    L = _lib()
    torch.manual_seed(0)
    M, vocab, ld = 5, 128256, 128256 + 64
    stream = torch.cuda.current_stream().cuda_stream
    ref = torch.zeros(M, dtype=torch.int32, device="cuda")
    got = torch.full((M,), -1, dtype=torch.int32, device="cuda")
    for call in range(2):                         # the shared array keeps state: the second call must be clean too
        logits = torch.randn(M, ld, device="cuda")
        logits[1, 777] = 50.0 + call
        logits[3, :vocab] = 0.0                   # ties -> lowest index
        logits[:, vocab:] = 1e9                   # padding columns beyond `vocab` must be ignored
        K.argmax(logits, M, vocab, ld, ref)       # existing entry point (caller-provided scratch)
        got.fill_(-1)
        torch.cuda.synchronize()
        rc = L.dl_argmax_noalloc(logits.data_ptr(), M, vocab, ld, got.data_ptr(), stream)
        assert rc == 0, L.dl_last_error().decode()
        torch.cuda.synchronize()
        assert got.tolist() == ref.tolist()
        assert got[1].item() == 777 and got[3].item() == 0
        assert got.tolist() == logits[:, :vocab].argmax(dim=-1).tolist()


def test_noalloc_argmax_rejects_sizes_beyond_its_scratch():
    L = _lib()
    x = torch.zeros(65, 32, device="cuda")
    out = torch.zeros(65, dtype=torch.int32, device="cuda")
    rc = L.dl_argmax_noalloc(x.data_ptr(), 65, 32, 32, out.data_ptr(), None)
    assert rc != 0 and b"64" in L.dl_last_error()
    assert L.dl_argmax_noalloc(x.data_ptr(), 0, 32, 32, out.data_ptr(), None) != 0
    assert L.dl_argmax_noalloc(x.data_ptr(), 64, 32, 32, out.data_ptr(), None) == 0     # the maximum itself works
    torch.cuda.synchronize()
    assert out[:64].tolist() == [0] * 64
