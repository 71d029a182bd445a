"""Q40 / Q80 block codecs (NumPy, host side) -- the bit-exact definition of the formats.

Formats (reference: src/nn/nn-quants.hpp:64-72, src/nn/nn-quants.cpp:67-246,
converter/writer.py:29-74):

* Q40 block = 18 bytes: fp16 scale ``d`` + 16 bytes.  Byte ``j`` holds element ``j`` in its low
  nibble and element ``j+16`` in its high nibble; value = ``(nibble - 8) * d``;
  ``d = x_maxabs_signed / -8``; nibble = ``min(15, trunc(x/d + 8.5))``.
* Q80 block = 34 bytes: fp16 scale ``d = amax/127`` + 32 int8; ``q = round_half_even(x * (1/d))``
  where ``1/d`` is computed from the *f32* ``d`` (nn-quants.cpp:126-128; AVX2 rounding mode
  ``_MM_FROUND_TO_NEAREST_INT`` :139, identical to ``np.round`` in converter/writer.py:67).

Blocks run along the input (n) axis of a weight row.  These NumPy routines are the oracle for the
C++ (csrc/host/quants.cpp) and CUDA (csrc/cuda/quant.cu) codecs.
"""
from __future__ import annotations

import numpy as np

QK = 32                    # elements per block (Q40 and Q80)
Q40_BLOCK_BYTES = 18
Q80_BLOCK_BYTES = 34

F32, F16, Q40, Q80 = 0, 1, 2, 3
FLOAT_TYPE_NAMES = {"f32": F32, "f16": F16, "q40": Q40, "q80": Q80}
FLOAT_TYPE_STR = {v: k for k, v in FLOAT_TYPE_NAMES.items()}


def parse_float_type(name: str) -> int:
    try:
        return FLOAT_TYPE_NAMES[name.lower()]
    except KeyError:
        raise ValueError(f"{name} is not a supported float type (f32, f16, q40, q80)")


def tensor_bytes(float_type: int, n_elements: int) -> int:
    """Serialized byte size of ``n_elements`` values of ``float_type``."""
    if float_type == F32:
        return 4 * n_elements
    if float_type == F16:
        return 2 * n_elements
    if float_type == Q40:
        assert n_elements % QK == 0
        return n_elements // QK * Q40_BLOCK_BYTES
    if float_type == Q80:
        assert n_elements % QK == 0
        return n_elements // QK * Q80_BLOCK_BYTES
    raise ValueError(f"unknown float type {float_type}")


# ----------------------------------------------------------------------------------------------
# Q40
# ----------------------------------------------------------------------------------------------
def quantize_q40(x: np.ndarray) -> np.ndarray:
    """f32 array (size % 32 == 0) -> uint8 array of Q40 blocks, shape [nBlocks, 18]."""
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, QK)
    gmax = x.max(axis=1)
    gmin = x.min(axis=1)
    deltas = (np.where(-gmin > gmax, gmin, gmax) / np.float32(-8.0)).astype(np.float32)
    with np.errstate(divide="ignore"):
        ids = np.where(deltas != 0, np.float32(1.0) / deltas, np.float32(0.0)).astype(np.float32)
    q = x * ids[:, None] + np.float32(8.5)
    q = np.clip(q, 0, 15).astype(np.int32)          # trunc toward zero == floor for q >= 0
    lo = q[:, :16] & 0xF
    hi = (q[:, 16:] & 0xF) << 4
    out = np.empty((x.shape[0], Q40_BLOCK_BYTES), dtype=np.uint8)
    out[:, 0:2] = deltas.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:] = (lo | hi).astype(np.uint8)
    return out


def dequantize_q40(blocks: np.ndarray) -> np.ndarray:
    """uint8 [nBlocks, 18] (or flat) -> f32 [nBlocks * 32]."""
    b = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, Q40_BLOCK_BYTES)
    d = b[:, 0:2].copy().view(np.float16).astype(np.float32).reshape(-1)
    qs = b[:, 2:]
    lo = (qs & 0x0F).astype(np.int32) - 8
    hi = (qs >> 4).astype(np.int32) - 8
    out = np.empty((b.shape[0], QK), dtype=np.float32)
    out[:, :16] = lo * d[:, None]
    out[:, 16:] = hi * d[:, None]
    return out.reshape(-1)


def q40_split(blocks: np.ndarray):
    """uint8 [..., 18] -> (scales fp16 [...], nibble bytes uint8 [..., 16])."""
    b = np.ascontiguousarray(blocks, dtype=np.uint8)
    b = b.reshape(-1, Q40_BLOCK_BYTES)
    d = b[:, 0:2].copy().view(np.float16).reshape(-1)
    return d, np.ascontiguousarray(b[:, 2:])


# ----------------------------------------------------------------------------------------------
# Q80
# ----------------------------------------------------------------------------------------------
def quantize_q80(x: np.ndarray) -> np.ndarray:
    """f32 array (size % 32 == 0) -> uint8 array of Q80 blocks, shape [nBlocks, 34]."""
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, QK)
    amax = np.abs(x).max(axis=1)
    d = (amax / np.float32(127.0)).astype(np.float32)
    with np.errstate(divide="ignore"):
        ids = np.where(d != 0, np.float32(1.0) / d, np.float32(0.0)).astype(np.float32)
    q = np.round(x * ids[:, None]).astype(np.int8)   # half-to-even
    out = np.empty((x.shape[0], Q80_BLOCK_BYTES), dtype=np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:] = q.view(np.uint8)
    return out


def dequantize_q80(blocks: np.ndarray) -> np.ndarray:
    b = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, Q80_BLOCK_BYTES)
    d = b[:, 0:2].copy().view(np.float16).astype(np.float32).reshape(-1)
    q = b[:, 2:].view(np.int8).astype(np.float32)
    return (q * d[:, None]).reshape(-1)


def q80_roundtrip(x: np.ndarray) -> np.ndarray:
    """quantize->dequantize along the last axis (what the reference's F32->Q80 cast loses)."""
    shape = x.shape
    return dequantize_q80(quantize_q80(x)).reshape(shape)


# ----------------------------------------------------------------------------------------------
# generic serializer used by the .m writer
# ----------------------------------------------------------------------------------------------
def encode_tensor(x: np.ndarray, float_type: int) -> bytes:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
    if float_type == F32:
        return x.tobytes()
    if float_type == F16:
        return x.astype(np.float16).tobytes()
    if float_type == Q40:
        return quantize_q40(x).tobytes()
    if float_type == Q80:
        return quantize_q80(x).tobytes()
    raise ValueError(f"unknown float type {float_type}")


def decode_tensor(buf, float_type: int, n_elements: int) -> np.ndarray:
    raw = np.frombuffer(buf, dtype=np.uint8, count=tensor_bytes(float_type, n_elements))
    if float_type == F32:
        return raw.view(np.float32).copy()
    if float_type == F16:
        return raw.view(np.float16).astype(np.float32)
    if float_type == Q40:
        return dequantize_q40(raw)
    if float_type == Q80:
        return dequantize_q80(raw)
    raise ValueError(f"unknown float type {float_type}")
