"""Meta `consolidated.*.pth` checkpoint -> `.m` model file.

    python -m dllama_b200.converter.convert_llama <modelPath> <q40|q80|f16|f32>

Behavioural parity with converter/convert-llama.py:10-98: ``params.json`` gives the header
(``max_seq_len`` and a valid ``vocab_size`` are required), ``hidden_dim`` is derived from the w1 shard
shape times the shard count, shards are concatenated along axis 1 for ``tok_embeddings``, ``wo`` and
``w2`` and along axis 0 otherwise, norms and the embedding stay F32, and tensors are processed in
chunks so that only a bounded set is resident at a time.  (Meta checkpoints already use the
interleaved-pair RoPE layout, so no permutation is applied.)
"""
from __future__ import annotations

import json
import os
import sys
from pathlib import Path

from ..formats import mfile
from ..ops import quant

LAYER_CHUNK_SIZE = 48


def layer_names(n_layers: int):
    names = ["tok_embeddings.weight"]
    for l in range(n_layers):
        p = f"layers.{l}."
        names += [p + "attention.wq.weight", p + "attention.wk.weight", p + "attention.wv.weight",
                  p + "attention.wo.weight", p + "feed_forward.w1.weight", p + "feed_forward.w2.weight",
                  p + "feed_forward.w3.weight", p + "attention_norm.weight", p + "ffn_norm.weight"]
    return names + ["norm.weight", "output.weight"]


def convert(model_path: str, out_path: str, weight_type: int, log=print) -> None:
    import torch
    with open(os.path.join(model_path, "params.json")) as f:
        params = json.load(f)
    if params.get("vocab_size", 0) < 1:
        raise ValueError("vocab_size is invalid, please update params.json file")
    if params.get("max_seq_len") is None:
        raise ValueError("max_seq_len is required, please update params.json file")
    shards = sorted(Path(model_path).glob("consolidated.*.pth"))
    if not shards:
        raise ValueError("no consolidated.*.pth files found")
    names = layer_names(params["n_layers"])
    header = dict(version=0, arch_type=mfile.ARCH_LLAMA, dim=params["dim"], n_layers=params["n_layers"],
                  n_heads=params["n_heads"], n_kv_heads=params.get("n_kv_heads") or params["n_heads"],
                  n_experts=0, n_active_experts=0, vocab_size=params["vocab_size"],
                  max_seq_len=params["max_seq_len"], hidden_act=mfile.HIDDEN_ACT_SILU,
                  weights_float_type=weight_type)
    if "rope_theta" in params:
        header["rope_theta"] = int(params["rope_theta"])
    wrote_header = False
    with open(out_path, "wb") as out:
        for c0 in range(0, len(names), LAYER_CHUNK_SIZE):
            chunk = names[c0:c0 + LAYER_CHUNK_SIZE]
            parts = {n: [] for n in chunk}
            log(f"💿 Chunking model {c0 // LAYER_CHUNK_SIZE + 1}/{-(-len(names) // LAYER_CHUNK_SIZE)}...")
            for sp in shards:
                model = torch.load(sp, map_location="cpu")
                for k in chunk:
                    if k in model:
                        parts[k].append(model[k])
                if not wrote_header:
                    header["hidden_dim"] = model["layers.0.feed_forward.w1.weight"].shape[0] * len(shards)
                    mfile.write_header(out, header)
                    wrote_header = True
                del model
            for name in chunk:
                axis1 = name == "tok_embeddings.weight" or name.endswith(".attention.wo.weight") or \
                    name.endswith(".feed_forward.w2.weight")
                f32 = name == "tok_embeddings.weight" or name.endswith("_norm.weight") or name == "norm.weight"
                ts = parts[name]
                if not ts:
                    raise KeyError(f"Layer {name} not found")
                t = ts[0] if len(ts) == 1 or ts[0].dim() == 1 else torch.cat(ts, dim=1 if axis1 else 0)
                log(f"🔶 Exporting {name} {list(t.shape)}...")
                out.write(quant.encode_tensor(t.to(torch.float32).numpy(), quant.F32 if f32 else weight_type))
            del parts


def main(argv=None) -> int:
    argv = sys.argv if argv is None else argv
    if len(argv) < 3:
        print("Usage: python -m dllama_b200.converter.convert_llama <modelPath> <targetFloatType>")
        return 1
    ft = quant.parse_float_type(argv[2])
    name = os.path.basename(os.path.normpath(argv[1])).lower()
    out = f"dllama_model_{name}_{argv[2]}.m"
    print(f"Target file: {out}")
    convert(argv[1], out, ft)
    print("Done!")
    return 0


if __name__ == "__main__":
    sys.exit(main())
