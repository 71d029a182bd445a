"""B2 baseline: the same model run the "library" way -- NCCL all-reduce + Q40 dequantisation +
cuBLAS matmuls (torch.matmul) + PyTorch elementwise ops.

This is the in-repo baseline BASELINE.json asks the fused kernels to beat ("a path that only calls
NCCL is the baseline, not the product").  It shares the weight slicing and the Q40T images with
``LlamaEngine`` (so both arms stream the same bytes) but:

* every matmul first dequantises its Q40T weight to bf16 (``dl_q40t_dequant`` + cast) and then
  calls cuBLAS through ``torch.matmul``;
* the partial sums of wo / w2 are combined with ``torch.distributed.all_reduce`` (NCCL);
* RMSNorm, RoPE, SiLU*mul, softmax attention are plain PyTorch ops on a dense (non-paged) KV cache;
* activations are Q80 round-tripped at the same points as the reference so the numerics are
  comparable with the oracle.

Single-sequence decode only (that is all the baseline is used for: bench.py --impl nccl_cublas).
"""
from __future__ import annotations

import math
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import kernels as K
from .llama import LlamaEngine
from .oracle import q80_rt


class NcclCublasEngine:
    def __init__(self, model_path: str, rank: int = 0, world: int = 1, max_seq_len: int = 2048,
                 cache_weights: bool = False, **_):
        # reuse the product engine only as a weight loader (per-rank slices, Q40T images)
        self.core = LlamaEngine(model_path, rank=rank, world=world, max_rows=8, n_slots=1,
                                max_seq_len=max_seq_len, use_graphs=False, use_pdl=False)
        c = self.core
        self.h, self.rank, self.world, self.device = c.h, rank, world, c.device
        self.seq_len = c.seq_len
        self.k_cache = [torch.zeros(self.seq_len, c.n_kv_loc, c.hs, device=c.device) for _ in range(c.h.n_layers)]
        self.v_cache = [torch.zeros(self.seq_len, c.n_kv_loc, c.hs, device=c.device) for _ in range(c.h.n_layers)]
        self.freqs = c.rope_freqs
        self.n_launch_est = 0
        # cache_weights: dequantise every matrix ONCE to bf16 (3.6x the HBM footprint of Q40) -- the stronger
        # variant of this baseline; the default re-dequantises on every call and keeps the Q40 footprint
        self.cache = {} if cache_weights else None

    # Q40T -> bf16 [d][n], every call (that is the point of this baseline)
    def _w(self, q40t: torch.Tensor, d: int, n: int) -> torch.Tensor:
        if self.cache is None:
            return K.q40t_dequant(q40t, d, n).to(torch.bfloat16)
        key = q40t.data_ptr()
        if key not in self.cache:
            self.cache[key] = K.q40t_dequant(q40t, d, n).to(torch.bfloat16)
        return self.cache[key]

    def _mm(self, x: torch.Tensor, q40t: torch.Tensor, d: int, n: int) -> torch.Tensor:
        return (q80_rt(x).to(torch.bfloat16) @ self._w(q40t, d, n).t()).float()

    def _rms(self, x, w):
        return w * (x * torch.rsqrt((x * x).mean(-1, keepdim=True) + self.h.norm_epsilon))

    def _rope(self, x, pos):
        T = x.shape[0]
        hs = self.core.hs
        xr = x.reshape(T, -1, hs // 2, 2)
        ang = pos.float()[:, None] * self.freqs[None, :]
        c, s = torch.cos(ang)[:, None, :], torch.sin(ang)[:, None, :]
        x0, x1 = xr[..., 0], xr[..., 1]
        return torch.stack([x0 * c - x1 * s, x0 * s + x1 * c], dim=-1).reshape(x.shape)

    @torch.no_grad()
    def forward(self, tokens: List[int], start_pos: int) -> torch.Tensor:
        c, h = self.core, self.h
        T = len(tokens)
        pos = torch.arange(start_pos, start_pos + T, device=self.device)
        x = c.embedding[torch.tensor(tokens, device=self.device)]
        G = c.n_heads_loc // c.n_kv_loc
        for l, W in enumerate(c.layers):
            y = self._rms(x, W["rms_att"])
            qkv = self._mm(y, W["qkv"], c.qkv_d, h.dim)
            q = self._rope(qkv[:, :c.q_loc], pos)
            k = self._rope(qkv[:, c.q_loc:c.q_loc + c.kv_loc], pos)
            v = qkv[:, c.q_loc + c.kv_loc:]
            self.k_cache[l][start_pos:start_pos + T] = k.reshape(T, c.n_kv_loc, c.hs)
            self.v_cache[l][start_pos:start_pos + T] = v.reshape(T, c.n_kv_loc, c.hs)
            kk = self.k_cache[l][:start_pos + T].repeat_interleave(G, dim=1)      # [S, heads, hs]
            vv = self.v_cache[l][:start_pos + T].repeat_interleave(G, dim=1)
            qh = q.reshape(T, c.n_heads_loc, c.hs)
            sc = torch.einsum("thd,shd->hts", qh, kk) / math.sqrt(c.hs)
            mask = torch.arange(start_pos + T, device=self.device)[None, :] > pos[:, None]
            sc = sc.masked_fill(mask[None], float("-inf"))
            att = torch.einsum("hts,shd->thd", torch.softmax(sc, dim=-1), vv).reshape(T, c.q_loc)
            part = q80_rt(self._mm(att, W["wo"], h.dim, c.q_loc))
            if self.world > 1:
                dist.all_reduce(part)
            x = x + part
            y = self._rms(x, W["rms_ffn"])
            gu = self._mm(y, W["w13"], 2 * c.hid_loc, h.dim)
            g, u = gu[:, 0::2], gu[:, 1::2]
            part = q80_rt(self._mm(torch.nn.functional.silu(g) * u, W["w2"], h.dim, c.hid_loc))
            if self.world > 1:
                dist.all_reduce(part)
            x = x + part
        y = self._rms(x, c.rms_final)
        lg = self._mm(y, c.wcls, c.vocab_loc, h.dim)
        if self.world > 1:
            out = [torch.empty_like(lg) for _ in range(self.world)]
            dist.all_gather(out, lg)
            lg = torch.cat(out, dim=-1)
        return lg

    # ---------------------------------------------------------------------------------------------------
    # The strongest form of this baseline: ONE CUDA graph per decode step, greedy pick on the device, no host
    # round trip per token (bench.py --impl nccl_cublas --b2-graph 1).  Same math as forward() for one row.
    # ---------------------------------------------------------------------------------------------------
    @torch.no_grad()
    def _step_static(self):
        c, h = self.core, self.h
        pos = self.s_pos                                            # [1] int64, device
        x = c.embedding[self.s_tok]                                 # [1, dim]
        G = c.n_heads_loc // c.n_kv_loc
        S = self.seq_len
        mask = torch.arange(S, device=self.device)[None, :] > pos[:, None]            # [1, S]
        for l, W in enumerate(c.layers):
            y = self._rms(x, W["rms_att"])
            qkv = self._mm(y, W["qkv"], c.qkv_d, h.dim)
            q = self._rope(qkv[:, :c.q_loc], pos)
            k = self._rope(qkv[:, c.q_loc:c.q_loc + c.kv_loc], pos)
            v = qkv[:, c.q_loc + c.kv_loc:]
            self.k_cache[l].index_copy_(0, pos, k.reshape(1, c.n_kv_loc, c.hs))
            self.v_cache[l].index_copy_(0, pos, v.reshape(1, c.n_kv_loc, c.hs))
            kk = self.k_cache[l].repeat_interleave(G, dim=1)          # whole cache, masked beyond pos
            vv = self.v_cache[l].repeat_interleave(G, dim=1)
            qh = q.reshape(1, c.n_heads_loc, c.hs)
            sc = torch.einsum("thd,shd->hts", qh, kk) / math.sqrt(c.hs)
            sc = sc.masked_fill(mask[None], float("-inf"))
            att = torch.einsum("hts,shd->thd", torch.softmax(sc, dim=-1), vv).reshape(1, c.q_loc)
            part = q80_rt(self._mm(att, W["wo"], h.dim, c.q_loc))
            if self.world > 1:
                dist.all_reduce(part)
            x = x + part
            y = self._rms(x, W["rms_ffn"])
            gu = self._mm(y, W["w13"], 2 * c.hid_loc, h.dim)
            g, u = gu[:, 0::2], gu[:, 1::2]
            part = q80_rt(self._mm(torch.nn.functional.silu(g) * u, W["w2"], h.dim, c.hid_loc))
            if self.world > 1:
                dist.all_reduce(part)
            x = x + part
        y = self._rms(x, c.rms_final)
        lg = self._mm(y, c.wcls, c.vocab_loc, h.dim)
        if self.world > 1:
            dist.all_gather_into_tensor(self.s_logits, lg.contiguous())
            lg = self.s_logits.reshape(self.world, -1).reshape(1, -1)
        self.s_tok.copy_(torch.argmax(lg, dim=-1))
        self.s_pos.add_(1)

    def start_graph_decode(self, token: int, pos: int):
        """Capture the decode step (weights must be cached: --b2-cache 1) and position it at (token, pos)."""
        self.s_tok = torch.tensor([token], dtype=torch.long, device=self.device)
        self.s_pos = torch.tensor([pos], dtype=torch.long, device=self.device)
        self.s_logits = torch.empty(self.world * self.core.vocab_loc, dtype=torch.float32, device=self.device)
        st = torch.cuda.Stream(device=self.device)
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            for _ in range(2):                                      # warm-up: cuBLAS handles, NCCL channels, weight cache
                self._step_static()
        torch.cuda.current_stream().wait_stream(st)
        torch.cuda.synchronize()
        self.s_tok.fill_(token)
        self.s_pos.fill_(pos)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=st):
            self._step_static()
        self.s_tok.fill_(token)
        self.s_pos.fill_(pos)
        torch.cuda.synchronize()

    def graph_step(self):
        self.graph.replay()

    def close(self):
        self.core.close()
