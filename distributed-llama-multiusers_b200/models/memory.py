"""Per-node memory estimate from the `.m` header alone (reference: printNodeRequiredMemory, src/nn/nn-core.cpp:162-176,
which sums the pipes and buffers of the node's graph).  Used by the CLI banner and to check a configuration against
the 180 GB of a B200 before anything is loaded (e.g. Llama-3.1-405B Q40 at TP=8)."""
from __future__ import annotations

from typing import Dict

from ..formats.mfile import LlmHeader
from ..ops import quant

B200_HBM_BYTES = 180 * 10 ** 9


def required_memory(h: LlmHeader, n_nodes: int = 1, max_rows: int = 32, n_slots: int = 4, seq_len: int = 0,
                    backend: str = "cuda") -> Dict[str, int]:
    """Bytes one node needs.  backend "cuda": Q40T tiles (same 4.5 bits/weight as the file, row count padded to 128),
    replicated f32 embedding, fp16 paged KV; backend "cpu": file-layout slices, embedding on the root only, f32 KV."""
    N = n_nodes
    seq = seq_len or h.seq_len
    wt = h.weight_type
    dim, hid, kv, vocab, L = h.dim, h.hidden_dim, h.kv_dim, h.vocab_size, h.n_layers

    def mat(rows: int, cols: int) -> int:
        if backend == "cuda" and wt == quant.Q40:
            rows = (rows + 127) // 128 * 128
        if backend == "cpu" and wt == quant.F16:
            return rows * cols * 4                      # f16 files are widened to f32 at load
        return rows * quant.tensor_bytes(wt, cols)

    per_layer = (mat(dim // N, dim) + 2 * mat(kv // N, dim) + mat(dim, dim // N) + 2 * mat(hid // N, dim)
                 + mat(dim, hid // N) + 2 * dim * 4)
    weights = L * per_layer + dim * 4 + mat(vocab // N, dim)
    embedding = vocab * dim * 4
    kv_elem = 2 if backend == "cuda" else 4
    kv_cache = 2 * L * n_slots * seq * (kv // N) * kv_elem
    # residual stream, normalised / quantised activations, q/k/v, hidden, logits, exchange area (N slices per sync)
    act = max_rows * (4 * dim * 4 + (dim + 2 * kv) // N * 4 + 3 * (hid // N) * 4 + vocab * 4 + 2 * N * dim * 8)
    out = {"weights": weights, "embedding": embedding, "kv_cache": kv_cache, "activations": act}
    # fp16 copy of the layer matrices for prompt-sized batches (LlamaEngine prefill_wh="resident": 2 bytes per weight,
    # rows padded to 128); only engines that can see >= 768-row batches build it, and "auto" declines it when it would
    # take more than a quarter of the free memory
    if backend == "cuda" and wt == quant.Q40 and max_rows >= 768:
        pad = lambda r: (r + 127) // 128 * 128
        out["fp16_weight_copy"] = 2 * L * (pad((dim + 2 * kv) // N) * dim + pad(dim) * (dim // N)
                                           + pad(2 * hid // N) * dim + pad(dim) * (hid // N))
    out["total"] = sum(out.values())
    return out


def describe(m: Dict[str, int]) -> str:
    mb = lambda b: f"{b / 2 ** 20:.0f} MB"
    extra = f", fp16 weight copy {mb(m['fp16_weight_copy'])}" if m.get("fp16_weight_copy") else ""
    return (f"📀 RequiredMemory: {mb(m['total'])} (weights {mb(m['weights'])}, embedding {mb(m['embedding'])}, "
            f"kv {mb(m['kv_cache'])}, activations {mb(m['activations'])}{extra})")
