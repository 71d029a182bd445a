"""Native CPU engine: the C++ host backend (csrc/host/cpu_backend.cpp) behind the engine surface of
``LlamaEngine`` -- batched rows with per-row position and KV slot, Q40 weights with Q80 activations / sync (or the
f32 path), and tensor parallelism across ``dllama worker`` processes over TCP.  Workers get their configuration
and weight slices from the root; they need no model file (reference: README.md:50, src/nn/nn-network.cpp:617-901).

It is what runs when there is no GPU (BASELINE.json's "root + 1 worker over localhost TCP, q80 buffer, CPU"
configuration).  ``CpuEngine`` (models/cpu_engine.py, pure PyTorch) stays as the numerics oracle.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..formats.mfile import read_header
from ..ops import quant
from ..text.native import host_lib

_sig_done = False


def _lib():
    global _sig_done
    L = host_lib()
    if not _sig_done:
        vp, i32, cp = C.c_void_p, C.c_int, C.c_char_p
        L.dl_cpu_root_new.restype = vp
        L.dl_cpu_root_new.argtypes = [cp, i32, C.POINTER(cp), C.POINTER(i32), i32, i32, i32, i32, i32, cp, i32]
        L.dl_cpu_forward.restype = i32
        L.dl_cpu_forward.argtypes = [vp, vp, vp, vp, i32, i32, vp]
        L.dl_cpu_last_error.restype = cp
        L.dl_cpu_last_error.argtypes = [vp]
        L.dl_cpu_stats.restype = None
        L.dl_cpu_stats.argtypes = [vp, C.POINTER(C.c_longlong), i32]
        L.dl_cpu_free.restype = None
        L.dl_cpu_free.argtypes = [vp]
        L.dl_cpu_worker_serve.restype = i32
        L.dl_cpu_worker_serve.argtypes = [i32, i32, cp, i32]
        L.dl_cpu_quantize_q80.argtypes = [vp, vp, i32, i32]
        L.dl_cpu_dequantize_q80.argtypes = [vp, vp, i32]
        L.dl_cpu_dequantize_q40.argtypes = [vp, vp, i32]
        L.dl_cpu_quantize_q40.argtypes = [vp, vp, i32]
        L.dl_cpu_f32_to_f16.restype = C.c_ushort
        L.dl_cpu_f32_to_f16.argtypes = [C.c_float]
        L.dl_cpu_f16_to_f32.restype = C.c_float
        L.dl_cpu_f16_to_f32.argtypes = [C.c_ushort]
        L.dl_cpu_matmul_q40_q80.argtypes = [vp, i32, i32, vp, i32, vp, i32]
        L.dl_cpu_has_avx2.restype = i32
        _sig_done = True
    return L


class _Slots:
    """Dense per-slot caches live in the backend; same surface as the GPU PageAllocator."""

    def __init__(self, n_slots: int):
        self.free_slots = list(range(n_slots - 1, -1, -1))

    def alloc_slot(self) -> int:
        if not self.free_slots:
            raise RuntimeError("no free sequence slot")
        return self.free_slots.pop()

    def free_slot(self, slot: int):
        self.free_slots.append(slot)

    def release_pages(self, slot: int):
        pass

    def ensure(self, slot: int, n_tokens: int):
        pass


class NativeCpuEngine:
    def __init__(self, model_path: str, workers: Sequence[Tuple[str, int]] = (), n_threads: int = 0, max_rows: int = 32,
                 n_slots: int = 4, max_seq_len: int = 0, sync_type: Optional[int] = None, **_unused):
        L = _lib()
        self.h = read_header(model_path, max_seq_len)
        if sync_type is None:
            sync_type = quant.Q80 if self.h.weight_type == quant.Q40 else quant.F32
        if sync_type not in (quant.Q80, quant.F32):
            raise ValueError("the sync (buffer) type must be q80 or f32")
        self.max_rows, self.n_slots = max_rows, n_slots
        self.seq_len = self.h.seq_len
        self.device = torch.device("cpu")
        self.rank, self.world = 0, 1 + len(workers)
        self.n_threads = n_threads if n_threads > 0 else max(1, min(16, (os.cpu_count() or 2) // max(1, self.world)))
        hosts = (C.c_char_p * max(1, len(workers)))(*[h.encode() for h, _ in workers])
        ports = (C.c_int * max(1, len(workers)))(*[p for _, p in workers])
        err = C.create_string_buffer(512)
        self._h = L.dl_cpu_root_new(model_path.encode(), len(workers), hosts, ports, self.n_threads, max_rows, n_slots,
                                    self.seq_len, 1 if sync_type == quant.Q80 else 0, err, 512)
        if not self._h:
            raise RuntimeError(err.value.decode("utf-8", "replace") or "cannot create the CPU engine")
        self.pages = _Slots(n_slots)
        self.logits = torch.zeros(max_rows, self.h.vocab_size)
        self._desc = np.zeros((3, max_rows), dtype=np.int32)

    def alloc_slot(self) -> int:
        return self.pages.alloc_slot()

    def free_slot(self, slot: int):
        self.pages.free_slot(slot)

    def ensure_capacity(self, slot: int, n_tokens: int):
        if n_tokens > self.seq_len:
            raise RuntimeError(f"sequence of {n_tokens} tokens exceeds max_seq_len")

    def forward(self, tokens: Sequence[int], positions: Sequence[int], slots: Sequence[int], logits_from: int = 0,
                greedy: bool = False) -> torch.Tensor:
        M = len(tokens)
        assert 1 <= M <= self.max_rows
        d = self._desc
        d[0, :M], d[1, :M], d[2, :M] = tokens, positions, slots
        rc = _lib().dl_cpu_forward(self._h, d[0].ctypes.data, d[1].ctypes.data, d[2].ctypes.data, M, logits_from,
                                   self.logits.data_ptr())
        if rc != 0:
            raise RuntimeError(_lib().dl_cpu_last_error(self._h).decode("utf-8", "replace"))
        return self.logits[:M - logits_from]

    def prefill(self, slot: int, tokens: Sequence[int], start_pos: int = 0, chunk: Optional[int] = None):
        chunk = chunk or self.max_rows
        out = None
        for i in range(0, len(tokens), chunk):
            part = list(tokens[i:i + chunk])
            pos = list(range(start_pos + i, start_pos + i + len(part)))
            out = self.forward(part, pos, [slot] * len(part), logits_from=len(part) - 1)
        return out

    def decode_step_greedy(self, tokens, positions, slots, logits_from: int = 0) -> List[int]:
        lg = self.forward(tokens, positions, slots, logits_from)
        return [int(t) for t in torch.argmax(lg, dim=-1)]

    def decode_greedy_host(self, tokens, positions, slots) -> List[int]:
        return self.decode_step_greedy(tokens, positions, slots, 0)

    def network_stats(self, reset: bool = True) -> dict:
        """Bytes over the sockets and time in node synchronisation since the last reset (reference:
        NnNetwork::getStats, the `Sync ms | Sent kB Recv kB` columns of `dllama inference`)."""
        out = (C.c_longlong * 3)()
        _lib().dl_cpu_stats(self._h, out, int(reset))
        return {"sent_bytes": int(out[0]), "recv_bytes": int(out[1]), "sync_us": int(out[2])}

    def close(self):
        if self._h:
            _lib().dl_cpu_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def serve_worker(port: int, n_threads: int = 0) -> None:
    """`dllama worker` without a GPU: serve one root until it stops or disconnects (blocking)."""
    err = C.create_string_buffer(512)
    rc = _lib().dl_cpu_worker_serve(port, n_threads, err, 512)
    if rc != 0:
        raise RuntimeError(err.value.decode("utf-8", "replace"))


# ---- codec helpers (tests; NumPy twins live in ops/quant.py) ---------------------------------------------------
def quantize_q80(x: np.ndarray, force_scalar: bool = False) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
    out = np.empty(x.size // 32 * 34, dtype=np.uint8)
    _lib().dl_cpu_quantize_q80(x.ctypes.data, out.ctypes.data, x.size, int(force_scalar))
    return out


def dequantize_q80(b: np.ndarray, n: int) -> np.ndarray:
    b = np.ascontiguousarray(b, dtype=np.uint8)
    y = np.empty(n, dtype=np.float32)
    _lib().dl_cpu_dequantize_q80(b.ctypes.data, y.ctypes.data, n)
    return y


def quantize_q40(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
    out = np.empty(x.size // 32 * 18, dtype=np.uint8)
    _lib().dl_cpu_quantize_q40(x.ctypes.data, out.ctypes.data, x.size)
    return out


def dequantize_q40(b: np.ndarray, n: int) -> np.ndarray:
    b = np.ascontiguousarray(b, dtype=np.uint8)
    y = np.empty(n, dtype=np.float32)
    _lib().dl_cpu_dequantize_q40(b.ctypes.data, y.ctypes.data, n)
    return y


def matmul_q40_q80(w_q40: np.ndarray, d: int, n: int, x_q80: np.ndarray, M: int, force_scalar: bool = False,
                   which: int = 0) -> np.ndarray:
    """which: 0 = the kernel the engine uses on this CPU (AVX-512 VNNI, AVX2 or scalar), 1 = scalar, 2 = AVX2."""
    y = np.empty((M, d), dtype=np.float32)
    _lib().dl_cpu_matmul_q40_q80(np.ascontiguousarray(w_q40).ctypes.data, d, n, np.ascontiguousarray(x_q80).ctypes.data, M,
                                 y.ctypes.data, 1 if force_scalar else which)
    return y
