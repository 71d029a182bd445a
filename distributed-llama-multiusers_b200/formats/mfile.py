"""`.m` model file: header parser/writer, tensor directory, random-init generator.

Layout (little endian; reference: converter/writer.py:109-145, src/llm.cpp:26-98,447-483,
src/llm.hpp:8-28):

    int32 magic = 0xA00ABCD
    int32 headerSize = 8 + 8 * nKv
    nKv x (int32 key, int32 value)
    tensors back to back, row-major [out][in], no padding:
        embed [vocab][dim]                       F32
        per layer: wq [dim][dim], wk [kvDim][dim], wv [kvDim][dim], wo [dim][dim],
                   w1/gate [hid][dim], w2/down [dim][hid], w3/up [hid][dim]   (weightType)
                   rms_att [dim] F32, rms_ffn [dim] F32
        rms_final [dim] F32
        wcls [vocab][dim]                        (weightType)

wq/wk rows are stored pre-permuted to interleaved-pair RoPE order by the converter
(converter/convert-hf.py:11-14).  All sizes are computed with Python ints (no 32-bit products;
cf. nn-core.cpp:109).
"""
from __future__ import annotations

import mmap
import os
import struct
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

from ..ops import quant

MAGIC = 0xA00ABCD
OLD_MAGICS = (0xABCD00, 0xABCD01)
ARCH_LLAMA = 0xABCD00

HEADER_KEYS = {
    "version": 0, "arch_type": 1, "dim": 2, "hidden_dim": 3, "n_layers": 4, "n_heads": 5,
    "n_kv_heads": 6, "n_experts": 7, "n_active_experts": 8, "vocab_size": 9, "max_seq_len": 10,
    "hidden_act": 11, "rope_theta": 12, "weights_float_type": 13, "rope_scaling_factor": 14,
    "rope_scaling_low_freq_factor": 15, "rope_scaling_high_freq_factory": 16,
    "rope_scaling_orig_max_seq_len": 17, "rope_type": 18,
}
KEY_NAMES = {v: k for k, v in HEADER_KEYS.items()}

HIDDEN_ACT_GELU, HIDDEN_ACT_SILU = 0, 1
ROPE_LLAMA, ROPE_FALCON, ROPE_LLAMA3_1 = 0, 1, 2


@dataclass
class LlmHeader:
    """Parsed `.m` header plus derived sizes (reference: LlmHeader, src/llm.hpp:39-67)."""
    version: int = 0
    arch_type: int = ARCH_LLAMA
    dim: int = 0
    hidden_dim: int = 0
    n_layers: int = 0
    n_heads: int = 0
    n_kv_heads: int = 0
    n_experts: int = 0
    n_active_experts: int = 0
    vocab_size: int = 0
    orig_seq_len: int = 0
    seq_len: int = 0
    hidden_act: int = HIDDEN_ACT_SILU
    rope_theta: float = 10000.0
    rope_type: int = ROPE_LLAMA
    rope_scaling_factor: float = 1.0
    rope_scaling_low_freq_factor: float = 0.0
    rope_scaling_high_freq_factor: float = 0.0
    rope_scaling_orig_max_seq_len: int = 0
    norm_epsilon: float = 1e-5
    weight_type: int = -1
    sync_type: int = quant.F32
    header_size: int = 0
    file_size: int = 0

    @property
    def head_size(self) -> int:
        return self.dim // self.n_heads

    @property
    def kv_dim(self) -> int:
        return self.dim * self.n_kv_heads // self.n_heads

    def describe(self) -> str:
        """Human-readable dump (reference: printLlmHeader, src/llm.cpp:100-124)."""
        rows = [
            ("Arch", "Llama"), ("HiddenAct", "Silu" if self.hidden_act == HIDDEN_ACT_SILU else "Gelu"),
            ("Dim", self.dim), ("KvDim", self.kv_dim), ("HiddenDim", self.hidden_dim),
            ("VocabSize", self.vocab_size), ("nLayers", self.n_layers), ("nHeads", self.n_heads),
            ("nKvHeads", self.n_kv_heads),
        ]
        if self.seq_len != self.orig_seq_len:
            rows.append(("OrigSeqLen", self.orig_seq_len))
        rows += [("SeqLen", self.seq_len), ("NormEpsilon", self.norm_epsilon),
                 ("RopeType", {0: "Llama", 1: "Falcon", 2: "Llama3.1"}.get(self.rope_type, "?")),
                 ("RopeTheta", int(self.rope_theta))]
        if self.rope_type == ROPE_LLAMA3_1:
            rows.append(("RopeScaling", "f=%.1f, l=%.1f, h=%.1f, o=%d" % (
                self.rope_scaling_factor, self.rope_scaling_low_freq_factor,
                self.rope_scaling_high_freq_factor, self.rope_scaling_orig_max_seq_len)))
        return "\n".join(f"💡 {k}: {v}" for k, v in rows)


def write_header(f, params: Dict[str, int]) -> int:
    """Serialize a header; returns bytes written (reference: writer.py:109-145)."""
    data = b""
    for key, value in params.items():
        if key not in HEADER_KEYS:
            raise KeyError(f"unknown header key {key}")
        data += struct.pack("<ii", HEADER_KEYS[key], int(value))
    head = struct.pack("<ii", MAGIC, 8 + len(data))
    f.write(head)
    f.write(data)
    return len(head) + len(data)


def read_header(path: str, max_seq_len: int = 0, sync_type: int = quant.F32) -> LlmHeader:
    """Parse the header of ``path`` (reference: loadLlmHeader, src/llm.cpp:26-98)."""
    h = LlmHeader()
    with open(path, "rb") as f:
        raw = f.read(8)
        if len(raw) != 8:
            raise ValueError("Cannot read magic value")
        magic, header_size = struct.unpack("<ii", raw)
        if magic in OLD_MAGICS:
            raise ValueError("Old model format is not supported")
        if magic != MAGIC:
            raise ValueError("Unsupported magic number")
        body = f.read(header_size - 8)
        if len(body) != header_size - 8:
            raise ValueError("Cannot read header values")
        f.seek(0, os.SEEK_END)
        h.file_size = f.tell()
    h.header_size = header_size
    kv = struct.unpack("<%di" % (len(body) // 4), body)
    for i in range(0, len(kv), 2):
        key, value = kv[i], kv[i + 1]
        name = KEY_NAMES.get(key)
        if name is None:
            raise ValueError("Unsupported header key")
        if name == "version": h.version = value
        elif name == "arch_type": h.arch_type = value
        elif name == "dim": h.dim = value
        elif name == "hidden_dim": h.hidden_dim = value
        elif name == "n_layers": h.n_layers = value
        elif name == "n_heads": h.n_heads = value
        elif name == "n_kv_heads": h.n_kv_heads = value
        elif name == "n_experts": h.n_experts = value
        elif name == "n_active_experts": h.n_active_experts = value
        elif name == "vocab_size": h.vocab_size = value
        elif name == "max_seq_len": h.seq_len = value
        elif name == "hidden_act": h.hidden_act = value
        elif name == "rope_theta": h.rope_theta = float(value)
        elif name == "weights_float_type": h.weight_type = value
        elif name == "rope_scaling_factor": h.rope_scaling_factor = float(value)
        elif name == "rope_scaling_low_freq_factor": h.rope_scaling_low_freq_factor = float(value)
        elif name == "rope_scaling_high_freq_factory": h.rope_scaling_high_freq_factor = float(value)
        elif name == "rope_scaling_orig_max_seq_len": h.rope_scaling_orig_max_seq_len = value
        elif name == "rope_type": h.rope_type = value
    if h.weight_type < 0:
        raise ValueError("Model does not specify weight type")
    if h.arch_type != ARCH_LLAMA:
        raise ValueError("Unsupported architecture")
    h.orig_seq_len = h.seq_len
    if max_seq_len > 0 and h.seq_len > max_seq_len:
        h.seq_len = max_seq_len
    h.sync_type = sync_type
    return h


@dataclass
class TensorEntry:
    name: str
    layer: int          # -1 for non-layer tensors
    rows: int           # output dim d
    cols: int           # input dim n
    float_type: int
    offset: int
    nbytes: int


def tensor_directory(h: LlmHeader) -> List[TensorEntry]:
    """Fixed tensor order of the file (reference: loadLlmNetWeight, src/llm.cpp:447-483)."""
    out: List[TensorEntry] = []
    off = h.header_size
    wt = h.weight_type

    def add(name, layer, rows, cols, ft):
        nonlocal off
        nb = rows * quant.tensor_bytes(ft, cols)
        out.append(TensorEntry(name, layer, rows, cols, ft, off, nb))
        off += nb

    add("embedding", -1, h.vocab_size, h.dim, quant.F32)
    for l in range(h.n_layers):
        add("wq", l, h.dim, h.dim, wt)
        add("wk", l, h.kv_dim, h.dim, wt)
        add("wv", l, h.kv_dim, h.dim, wt)
        add("wo", l, h.dim, h.dim, wt)
        add("w1", l, h.hidden_dim, h.dim, wt)
        add("w2", l, h.dim, h.hidden_dim, wt)
        add("w3", l, h.hidden_dim, h.dim, wt)
        add("rms_att", l, 1, h.dim, quant.F32)
        add("rms_ffn", l, 1, h.dim, quant.F32)
    add("rms_final", -1, 1, h.dim, quant.F32)
    add("wcls", -1, h.vocab_size, h.dim, wt)
    return out


def expected_file_size(h: LlmHeader) -> int:
    d = tensor_directory(h)
    return d[-1].offset + d[-1].nbytes


class ModelFile:
    """Read-only mmap view of a `.m` file (reference: src/mmap.hpp:35-81 + the walker above)."""

    def __init__(self, path: str, max_seq_len: int = 0, sync_type: int = quant.F32):
        self.path = path
        self.header = read_header(path, max_seq_len, sync_type)
        self.entries = tensor_directory(self.header)
        self.index: Dict[Tuple[str, int], TensorEntry] = {(e.name, e.layer): e for e in self.entries}
        want = expected_file_size(self.header)
        if want != self.header.file_size:
            raise ValueError(f"Missing bytes in weight file: expected {want}, file has "
                             f"{self.header.file_size}")
        self._f = open(path, "rb")
        self._mm = mmap.mmap(self._f.fileno(), 0, access=mmap.ACCESS_READ)
        self.data = np.frombuffer(self._mm, dtype=np.uint8)

    def close(self):
        self.data = None
        try:
            self._mm.close()
        except BufferError:
            pass
        self._f.close()

    def entry(self, name: str, layer: int = -1) -> TensorEntry:
        return self.index[(name, layer)]

    def raw(self, name: str, layer: int = -1) -> np.ndarray:
        e = self.entry(name, layer)
        return self.data[e.offset:e.offset + e.nbytes]

    def raw_rows(self, name: str, layer: int, row0: int, nrows: int) -> np.ndarray:
        """Bytes of rows [row0, row0+nrows) of a matrix (zero-copy view)."""
        e = self.entry(name, layer)
        rb = e.nbytes // e.rows
        return self.data[e.offset + row0 * rb: e.offset + (row0 + nrows) * rb]

    def f32(self, name: str, layer: int = -1) -> np.ndarray:
        e = self.entry(name, layer)
        return quant.decode_tensor(self.raw(name, layer), e.float_type, e.rows * e.cols).reshape(
            e.rows, e.cols)


# ----------------------------------------------------------------------------------------------
# model zoo shapes + random-init writer (there is no network: weights are synthetic)
# ----------------------------------------------------------------------------------------------
MODEL_SHAPES = {
    # name: dim, hidden, layers, heads, kv_heads, vocab, seq_len, rope (factor, low, high, orig)
    "tiny":        dict(dim=256, hidden_dim=512, n_layers=2, n_heads=8, n_kv_heads=4, vocab_size=1024,
                        max_seq_len=512, rope_theta=10000, rope=None),
    "tiny-tp8":    dict(dim=1024, hidden_dim=2048, n_layers=2, n_heads=16, n_kv_heads=8,
                        vocab_size=2048, max_seq_len=512, rope_theta=500000, rope=(32, 1, 4, 8192)),
    "llama3.2-1b": dict(dim=2048, hidden_dim=8192, n_layers=16, n_heads=32, n_kv_heads=8,
                        vocab_size=128256, max_seq_len=131072, rope_theta=500000, rope=(32, 1, 4, 8192)),
    "llama3.1-8b": dict(dim=4096, hidden_dim=14336, n_layers=32, n_heads=32, n_kv_heads=8,
                        vocab_size=128256, max_seq_len=131072, rope_theta=500000, rope=(8, 1, 4, 8192)),
    "llama3.3-70b": dict(dim=8192, hidden_dim=28672, n_layers=80, n_heads=64, n_kv_heads=8,
                         vocab_size=128256, max_seq_len=131072, rope_theta=500000, rope=(8, 1, 4, 8192)),
    "llama3.1-405b": dict(dim=16384, hidden_dim=53248, n_layers=126, n_heads=128, n_kv_heads=8,
                          vocab_size=128256, max_seq_len=131072, rope_theta=500000, rope=(8, 1, 4, 8192)),
}


def header_params(shape: Dict, weight_type: int, n_layers: Optional[int] = None) -> Dict[str, int]:
    p = dict(version=0, arch_type=ARCH_LLAMA, dim=shape["dim"], hidden_dim=shape["hidden_dim"],
             n_layers=n_layers or shape["n_layers"], n_heads=shape["n_heads"],
             n_kv_heads=shape["n_kv_heads"], n_experts=0, n_active_experts=0,
             vocab_size=shape["vocab_size"], max_seq_len=shape["max_seq_len"],
             hidden_act=HIDDEN_ACT_SILU, rope_theta=int(shape["rope_theta"]),
             weights_float_type=weight_type)
    if shape.get("rope"):
        f, lo, hi, orig = shape["rope"]
        p.update(rope_scaling_factor=int(f), rope_scaling_low_freq_factor=int(lo),
                 rope_scaling_high_freq_factory=int(hi), rope_scaling_orig_max_seq_len=int(orig),
                 rope_type=ROPE_LLAMA3_1)
    return p


def _random_q40_bytes(rng: np.random.Generator, n_blocks: int, scale: float) -> np.ndarray:
    """Random Q40 blocks drawn directly in the quantized domain (fast path for multi-GB files)."""
    out = np.empty((n_blocks, quant.Q40_BLOCK_BYTES), dtype=np.uint8)
    d = (rng.random(n_blocks, dtype=np.float32) * 0.5 + 0.75) * np.float32(scale / 8.0)
    d *= np.where(rng.integers(0, 2, n_blocks, dtype=np.int8) == 0, np.float32(1), np.float32(-1))
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:] = rng.integers(0, 256, (n_blocks, 16), dtype=np.uint8)
    return out


def write_random_model(path: str, shape_name_or_dict, weight_type: int = quant.Q40, seed: int = 1,
                       n_layers: Optional[int] = None, std: float = 0.02, exact: bool = False,
                       chunk_rows: int = 4096, native: Optional[bool] = None) -> LlmHeader:
    """Write a random-init `.m` of the given architecture.

    ``exact=True`` draws N(0,std) f32 weights and runs the real quantizer (small models / tests);
    otherwise Q40 blocks are drawn directly in the quantized domain (uniform nibbles, scales
    ~ std/8), which is statistically equivalent for benchmarking and ~100x faster for the 6 GB
    8B file.  ``native=True`` (default for files above 16 GB, or DLLAMA_FAST_WRITER=1) fills the file from several
    threads in C++ (csrc/host/model_writer.cpp; same recipe, different random stream): the 238 GB Llama-3.1-405B
    file takes a minute instead of twenty.
    """
    shape = MODEL_SHAPES[shape_name_or_dict] if isinstance(shape_name_or_dict, str) else shape_name_or_dict
    params = header_params(shape, weight_type, n_layers)
    if native is None:
        n_par = params["n_layers"] * (2 * params["dim"] ** 2 + 3 * params["dim"] * params["hidden_dim"])
        native = os.environ.get("DLLAMA_FAST_WRITER") == "1" or n_par * 0.5625 > 16e9
    if native and weight_type == quant.Q40 and not exact:
        import ctypes as C
        from ..text.native import host_lib
        L = host_lib()
        L.dl_write_random_model_q40.restype = C.c_longlong
        L.dl_write_random_model_q40.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.c_int, C.c_ulonglong, C.c_float, C.c_int]
        kv = []
        for key, value in params.items():
            kv += [HEADER_KEYS[key], int(value)]
        arr = (C.c_int * len(kv))(*kv)
        rc = L.dl_write_random_model_q40(path.encode(), arr, len(kv) // 2, seed, std, 0)
        if rc < 0:
            raise OSError(f"native model writer failed ({rc}) for {path}")
        return read_header(path)
    rng = np.random.default_rng(seed)
    dim, hid = params["dim"], params["hidden_dim"]
    kv_dim = dim * params["n_kv_heads"] // params["n_heads"]
    vocab = params["vocab_size"]
    tmp = path + ".tmp%d" % os.getpid()

    def write_matrix(f, rows, cols, ft, s):
        for r0 in range(0, rows, chunk_rows):
            r = min(chunk_rows, rows - r0)
            if ft == quant.Q40 and not exact:
                f.write(_random_q40_bytes(rng, r * cols // quant.QK, s * 3.0).tobytes())
            else:
                x = rng.standard_normal((r, cols), dtype=np.float32) * np.float32(s)
                f.write(quant.encode_tensor(x, ft))

    with open(tmp, "wb") as f:
        write_header(f, params)
        write_matrix(f, vocab, dim, quant.F32, 1.0 if not exact else std * 50)
        for _ in range(params["n_layers"]):
            write_matrix(f, dim, dim, weight_type, std)         # wq
            write_matrix(f, kv_dim, dim, weight_type, std)      # wk
            write_matrix(f, kv_dim, dim, weight_type, std)      # wv
            write_matrix(f, dim, dim, weight_type, std)         # wo
            write_matrix(f, hid, dim, weight_type, std)         # w1
            write_matrix(f, dim, hid, weight_type, std)         # w2
            write_matrix(f, hid, dim, weight_type, std)         # w3
            f.write((1.0 + 0.1 * rng.standard_normal(dim, dtype=np.float32)).astype(np.float32).tobytes())
            f.write((1.0 + 0.1 * rng.standard_normal(dim, dtype=np.float32)).astype(np.float32).tobytes())
        f.write((1.0 + 0.1 * rng.standard_normal(dim, dtype=np.float32)).astype(np.float32).tobytes())
        write_matrix(f, vocab, dim, weight_type, std)           # wcls
    os.replace(tmp, path)
    return read_header(path)
