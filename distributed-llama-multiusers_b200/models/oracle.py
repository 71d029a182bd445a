"""Pure-PyTorch CPU oracle of the reference's Llama forward graph.

It reproduces *where* the reference loses precision (reference
src/llm.cpp:126-438 and src/nn/nn-cpu-ops.cpp): Q40 weights, every matmul input cast to Q80
(att-4, att-13, ff-4, ff-9, end-4), and -- when ``n_nodes > 1`` -- every per-node partial sum of
wo/w2 cast to Q80 before the merge-add (att-15, ff-11).  Residual stream, norms, RoPE, KV cache,
attention, SiLU*mul and logits are f32.  It is the numerics reference for every CUDA kernel test
and for CPU-only CI; it is deliberately simple and slow.
"""
from __future__ import annotations

import math
from typing import List, Optional

import numpy as np
import torch

from ..formats.mfile import LlmHeader, ModelFile, ROPE_LLAMA3_1, HIDDEN_ACT_SILU
from ..ops import quant


def q80_rt(x: torch.Tensor) -> torch.Tensor:
    """F32 -> Q80 -> F32 along the last axis, torch version (half-even rounding)."""
    shp = x.shape
    g = x.reshape(-1, quant.QK).to(torch.float32)
    amax = g.abs().amax(dim=1)
    d = amax / 127.0
    inv = torch.where(d != 0, 1.0 / d, torch.zeros_like(d))
    q = torch.round(g * inv[:, None])
    d16 = d.to(torch.float16).to(torch.float32)
    return (q * d16[:, None]).reshape(shp)


def rope_frequencies(h: LlmHeader) -> torch.Tensor:
    """Per-pair rotation frequency, f32 [head_size/2] (reference: nn-core.cpp:307-340)."""
    hs = h.head_size
    out = np.empty(hs // 2, dtype=np.float32)
    for j in range(hs // 2):
        head_dim = 2 * j
        freq = np.float32(1.0) / np.float32(math.pow(np.float32(h.rope_theta), np.float32(head_dim) / np.float32(hs)))
        if h.rope_type == ROPE_LLAMA3_1 and h.rope_scaling_factor != 1.0:
            wave_len = np.float32(2.0 * math.pi) / freq
            high_wl = np.float32(h.rope_scaling_orig_max_seq_len / h.rope_scaling_high_freq_factor)
            low_wl = np.float32(h.rope_scaling_orig_max_seq_len / h.rope_scaling_low_freq_factor)
            if wave_len < high_wl:
                pass
            elif wave_len > low_wl:
                freq = freq / np.float32(h.rope_scaling_factor)
            else:
                smooth = (np.float32(h.rope_scaling_orig_max_seq_len) / wave_len
                          - np.float32(h.rope_scaling_low_freq_factor)) / np.float32(
                    h.rope_scaling_high_freq_factor - h.rope_scaling_low_freq_factor)
                freq = (1 - smooth) * freq / np.float32(h.rope_scaling_factor) + smooth * freq
        out[j] = freq
    return torch.from_numpy(out)


def apply_rope(x: torch.Tensor, pos: torch.Tensor, freqs: torch.Tensor, head_size: int) -> torch.Tensor:
    """Rotate adjacent pairs (2i, 2i+1) (reference: ropeLlamaForward_F32_F32, nn-cpu-ops.cpp:1090-1120)."""
    T = x.shape[0]
    xr = x.reshape(T, -1, head_size // 2, 2)
    ang = pos.to(torch.float32)[:, None] * freqs[None, :]          # [T, hs/2]
    c, s = torch.cos(ang)[:, None, :], torch.sin(ang)[:, None, :]
    x0, x1 = xr[..., 0], xr[..., 1]
    out = torch.stack([x0 * c - x1 * s, x0 * s + x1 * c], dim=-1)
    return out.reshape(x.shape)


class OracleLlama:
    """Dequantized f32 copy of a `.m` model with the reference's quantization points."""

    def __init__(self, mf: ModelFile, n_nodes: int = 1, sync_q80: Optional[bool] = None,
                 max_seq_len: Optional[int] = None):
        h = mf.header
        self.h = h
        self.n_nodes = n_nodes
        # q40 weights imply q80 activations/sync (reference: app.cpp:239-240)
        self.q80 = (h.weight_type == quant.Q40) if sync_q80 is None else sync_q80
        self.seq_len = max_seq_len or min(h.seq_len, 2048)
        t = lambda name, l=-1: torch.from_numpy(mf.f32(name, l))
        self.embedding = t("embedding")
        self.layers = []
        for l in range(h.n_layers):
            self.layers.append({k: t(k, l) for k in
                                ("wq", "wk", "wv", "wo", "w1", "w2", "w3", "rms_att", "rms_ffn")})
        self.rms_final = t("rms_final")
        self.wcls = t("wcls")
        self.freqs = rope_frequencies(h)
        self.reset()

    def reset(self):
        h = self.h
        self.k_cache = [torch.zeros(self.seq_len, h.kv_dim) for _ in range(h.n_layers)]
        self.v_cache = [torch.zeros(self.seq_len, h.kv_dim) for _ in range(h.n_layers)]

    def _cast(self, x):
        return q80_rt(x) if self.q80 else x

    def _rmsnorm(self, x, w):
        inv = 1.0 / torch.sqrt((x * x).mean(dim=-1, keepdim=True) + self.h.norm_epsilon)
        return w.reshape(1, -1) * (inv * x)

    def _col_parallel_sum(self, xin: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
        """wo / w2: input-split matmul; per-node partial sums are Q80-cast before the merge."""
        N = self.n_nodes
        xin = self._cast(xin)
        if N == 1:
            # the reference still casts the (single) partial to the sync type (llm.cpp:308-313)
            return self._cast(xin @ w.t())
        n = w.shape[1] // N
        acc = None
        for r in range(N):
            part = xin[:, r * n:(r + 1) * n] @ w[:, r * n:(r + 1) * n].t()
            part = self._cast(part)
            acc = part if acc is None else acc + part
        return acc

    @torch.no_grad()
    def forward(self, tokens: List[int], start_pos: int) -> torch.Tensor:
        """Run ``tokens`` at positions start_pos.. ; returns logits [T, vocab] (f32)."""
        h = self.h
        T = len(tokens)
        pos = torch.arange(start_pos, start_pos + T)
        x = self.embedding[torch.tensor(tokens, dtype=torch.long)].clone()
        hs, n_rep = h.head_size, h.n_heads // h.n_kv_heads
        for l, L in enumerate(self.layers):
            y = self._cast(self._rmsnorm(x, L["rms_att"]))
            q = apply_rope(y @ L["wq"].t(), pos, self.freqs, hs)
            k = apply_rope(y @ L["wk"].t(), pos, self.freqs, hs)
            v = y @ L["wv"].t()
            self.k_cache[l][start_pos:start_pos + T] = k
            self.v_cache[l][start_pos:start_pos + T] = v
            att_out = torch.empty(T, h.dim)
            for t in range(T):
                p = start_pos + t
                kk = self.k_cache[l][:p + 1].reshape(p + 1, h.n_kv_heads, hs)
                vv = self.v_cache[l][:p + 1].reshape(p + 1, h.n_kv_heads, hs)
                qq = q[t].reshape(h.n_heads, hs)
                kk = kk.repeat_interleave(n_rep, dim=1)
                vv = vv.repeat_interleave(n_rep, dim=1)
                sc = torch.einsum("hd,thd->ht", qq, kk) / math.sqrt(hs)
                a = torch.softmax(sc, dim=-1)
                att_out[t] = torch.einsum("ht,thd->hd", a, vv).reshape(-1)
            x = x + self._col_parallel_sum(att_out, L["wo"])
            y = self._cast(self._rmsnorm(x, L["rms_ffn"]))
            g = y @ L["w1"].t()
            u = y @ L["w3"].t()
            if h.hidden_act == HIDDEN_ACT_SILU:
                g = g / (1.0 + torch.exp(-g))
            else:
                g = 0.5 * g * (1.0 + torch.tanh(0.797884560802865 * (g + 0.044715 * g ** 3)))
            x = x + self._col_parallel_sum(g * u, L["w2"])
        y = self._cast(self._rmsnorm(x, self.rms_final))
        return y @ self.wcls.t()

    def generate_greedy(self, prompt: List[int], n_new: int) -> List[int]:
        self.reset()
        logits = self.forward(prompt, 0)
        out = []
        pos = len(prompt)
        tok = int(torch.argmax(logits[-1]))
        for _ in range(n_new):
            out.append(tok)
            logits = self.forward([tok], pos)
            pos += 1
            tok = int(torch.argmax(logits[-1]))
        return out
