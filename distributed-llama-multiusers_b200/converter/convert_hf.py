"""Hugging Face (safetensors) checkpoint -> `.m` model file.

    python -m dllama_b200.converter.convert_hf <sourceFolder> <q40|q80|f16|f32> <name>

Behavioural parity with the reference converter (converter/convert-hf.py): ``config.json`` is
mapped to the header (:152-195, all values int32 including rope_theta and the rope scaling
factors), safetensors shards are streamed one at a time (:31-43), q/k rows are permuted from HF's
half-split RoPE layout to the interleaved-pair layout the runtime rotates (:11-14,45-49), tensors
are written in runtime order (:51-89) and ``lm_head`` falls back to ``embed_tokens`` for tied
models (:88-89).  Accepts ``llama`` and ``mistral`` model types.  Mixture-of-experts checkpoints are
rejected up front: the runtime has no router/expert op (the reference plans those tensors,
:66-73, but cannot load them either).

The safetensors container is parsed directly (8-byte little-endian header length, JSON index, raw
little-endian tensor data), so only NumPy/PyTorch are needed.
"""
from __future__ import annotations

import json
import mmap
import os
import struct
import sys
from typing import Dict, List, Tuple

import numpy as np

from ..formats import mfile
from ..ops import quant

_DTYPES = {"F32": (np.float32, 4), "F16": (np.float16, 2), "BF16": (np.uint16, 2), "F64": (np.float64, 8)}


class SafetensorsFile:
    def __init__(self, path: str):
        self.path = path
        self._f = open(path, "rb")
        (n,) = struct.unpack("<Q", self._f.read(8))
        self.index: Dict[str, dict] = json.loads(self._f.read(n).decode("utf-8"))
        self.index.pop("__metadata__", None)
        self._base = 8 + n
        self._mm = mmap.mmap(self._f.fileno(), 0, access=mmap.ACCESS_READ)

    def keys(self) -> List[str]:
        return list(self.index.keys())

    def get(self, name: str) -> np.ndarray:
        e = self.index[name]
        dt, size = _DTYPES[e["dtype"]]
        a, b = e["data_offsets"]
        raw = np.frombuffer(self._mm, dtype=dt, count=(b - a) // size, offset=self._base + a)
        if e["dtype"] == "BF16":
            raw = (raw.astype(np.uint32) << 16).view(np.float32)
        return raw.astype(np.float32, copy=False).reshape(e["shape"])

    def close(self):
        self._mm.close()
        self._f.close()


def permute_rope_rows(w: np.ndarray, n_heads: int) -> np.ndarray:
    """HF stores each head's rotary dims as [first halves | second halves]; the runtime rotates adjacent
    pairs (2i, 2i+1).  Reorder the output rows of q/k accordingly."""
    rows = w.shape[0]
    return w.reshape(n_heads, 2, rows // n_heads // 2, *w.shape[1:]).swapaxes(1, 2).reshape(w.shape)


def load_config(folder: str, weight_type: int) -> Tuple[Dict[str, int], List[str]]:
    with open(os.path.join(folder, "config.json")) as f:
        cfg = json.load(f)
    if cfg.get("model_type") not in ("llama", "mistral"):
        raise ValueError(f"Unsupported arch type: {cfg.get('model_type')}")
    act = {"gelu": mfile.HIDDEN_ACT_GELU, "silu": mfile.HIDDEN_ACT_SILU}.get(cfg.get("hidden_act"))
    if act is None:
        raise ValueError(f"Unsupported hidden act: {cfg.get('hidden_act')}")
    if cfg.get("num_local_experts"):
        raise ValueError("mixture-of-experts checkpoints are not supported by the runtime")
    files = sorted(os.path.join(folder, f) for f in os.listdir(folder)
                   if f.endswith(".safetensors") and not f.startswith("."))
    if not files:
        raise ValueError("Not found any model file")
    p = dict(version=0, arch_type=mfile.ARCH_LLAMA, hidden_act=act, dim=cfg["hidden_size"],
             hidden_dim=cfg["intermediate_size"], n_layers=cfg["num_hidden_layers"],
             n_heads=cfg["num_attention_heads"], n_kv_heads=cfg.get("num_key_value_heads", cfg["num_attention_heads"]),
             weights_float_type=weight_type, max_seq_len=cfg["max_position_embeddings"], vocab_size=cfg["vocab_size"],
             n_experts=0, n_active_experts=0)
    if cfg.get("rope_theta") is not None:
        p["rope_theta"] = int(cfg["rope_theta"])
    rs = cfg.get("rope_scaling")
    if rs is not None:
        if rs.get("rope_type", rs.get("type")) != "llama3":
            raise ValueError(f"Unsupported rope type: {rs.get('rope_type')}")
        p.update(rope_scaling_factor=int(rs["factor"]), rope_scaling_low_freq_factor=int(rs["low_freq_factor"]),
                 rope_scaling_high_freq_factory=int(rs["high_freq_factor"]),
                 rope_scaling_orig_max_seq_len=int(rs["original_max_position_embeddings"]),
                 rope_type=mfile.ROPE_LLAMA3_1)
    return p, files


def convert(folder: str, weight_type: int, out_path: str, log=print) -> None:
    params, files = load_config(folder, weight_type)
    shards = [SafetensorsFile(f) for f in files]
    where = {k: s for s in shards for k in s.keys()}

    def fetch(*names) -> np.ndarray:
        for n in names:
            if n in where:
                return where[n].get(n)
        raise KeyError(f"Layer {names[0]} not found")

    wt, nh, nkv = weight_type, params["n_heads"], params["n_kv_heads"]
    with open(out_path, "wb") as out:
        mfile.write_header(out, params)

        def put(arr: np.ndarray, ft: int, name: str):
            log(f"🔶 Writing tensor {name} {list(arr.shape)}...")
            out.write(quant.encode_tensor(arr, ft))

        put(fetch("model.embed_tokens.weight"), quant.F32, "embed_tokens")
        for l in range(params["n_layers"]):
            pre = f"model.layers.{l}."
            put(permute_rope_rows(fetch(pre + "self_attn.q_proj.weight"), nh), wt, pre + "q_proj")
            put(permute_rope_rows(fetch(pre + "self_attn.k_proj.weight"), nkv), wt, pre + "k_proj")
            put(fetch(pre + "self_attn.v_proj.weight"), wt, pre + "v_proj")
            put(fetch(pre + "self_attn.o_proj.weight"), wt, pre + "o_proj")
            put(fetch(pre + "mlp.gate_proj.weight"), wt, pre + "gate_proj")
            put(fetch(pre + "mlp.down_proj.weight"), wt, pre + "down_proj")
            put(fetch(pre + "mlp.up_proj.weight"), wt, pre + "up_proj")
            put(fetch(pre + "input_layernorm.weight"), quant.F32, pre + "input_layernorm")
            put(fetch(pre + "post_attention_layernorm.weight"), quant.F32, pre + "post_attention_layernorm")
        put(fetch("model.norm.weight"), quant.F32, "model.norm")
        put(fetch("lm_head.weight", "model.embed_tokens.weight"), wt, "lm_head")
    for s in shards:
        s.close()


def main(argv=None) -> int:
    argv = sys.argv if argv is None else argv
    if len(argv) < 4:
        print("Usage: python -m dllama_b200.converter.convert_hf <sourceFolderPath> <weightsFloatType> <name>")
        return 1
    ft = quant.parse_float_type(argv[2])
    out = f"dllama_model_{argv[3]}_{argv[2]}.m"
    print(f"Output file: {out}")
    convert(argv[1], ft, out)
    print(f"✅ {out} created successfully")
    return 0


if __name__ == "__main__":
    sys.exit(main())
