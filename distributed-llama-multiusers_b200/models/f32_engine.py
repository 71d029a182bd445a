"""GPU engine for models whose weights are F32 / F16 in the `.m` file, run with an f32 sync buffer
(`--buffer-float-type f32`): the reference's "f32 / f32" path (matmul_F32_F32_F32, nn-cpu-ops.cpp:
182-220; merge-add of f32 slices :835-852; no Q80 cast anywhere).

Weights stay f32 in HBM.  On a GPU the whole forward pass runs on this repo's kernels: `gemm_f32_kernel` (plain f32
FMAs, weights streamed once per <= 8 activation rows), `reduce_norm_quant_kernel` (merge-add of the ranks' f32 partial
sums + residual + RMSNorm, f32 output), `rope_kv_f32` (RoPE + paged KV-cache write), `attention_kernel` over an f32
paged cache, `silu_mul_f32`, and for tensor parallelism `push_f32` (this rank's partial sums stored into every rank's
exchange buffer over NVLink + release flags: the reference's SYNC_NODE_SLICES of f32 slices) -- no cuBLAS, no NCCL on the
data path, no eager PyTorch ops.  The tuned, fused kernels of this repo are Q40 x Q80 (what the published models use);
this path is about running f32 models natively at all.  Without a GPU the same engine falls back to PyTorch CPU ops.
What it shares with ``LlamaEngine`` is the public surface -- slots, batched rows with per-row positions, tensor-parallel
slicing by rank -- so the scheduler, the HTTP API and the CLI run f32 models on GPUs unchanged.

Slicing follows the reference (llm.cpp:131-142): q/k/v/w1/w3/wcls by output rows, wo/w2 by input
columns, KV cache by kv head, residual stream replicated.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from ..formats.mfile import HIDDEN_ACT_SILU, ModelFile
from ..parallel.symm import validate_tp
from .oracle import rope_frequencies


class _Slots:
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


class F32Engine:
    def __init__(self, model_path: str, rank: int = 0, world: int = 1, group=None, max_rows: int = 32,
                 n_slots: int = 4, max_seq_len: int = 2048, device: Optional[str] = None, **_unused):
        self.mf = ModelFile(model_path, max_seq_len=max_seq_len)
        h = self.h = self.mf.header
        validate_tp(world, h.n_kv_heads, h.dim, h.hidden_dim, h.n_heads, h.n_heads, q40=False)      # vocab may be replicated
        self.rank, self.world, self.group = rank, world, group
        self.max_rows, self.n_slots = max_rows, n_slots
        self.seq_len = min(h.seq_len, max_seq_len)
        self.device = torch.device(device or (f"cuda:{torch.cuda.current_device()}" if torch.cuda.is_available() else "cpu"))
        if self.device.type == "cuda":
            torch.backends.cuda.matmul.allow_tf32 = False          # the reference accumulates in f32
        N, hs = world, h.head_size
        self.hs = hs
        self.n_heads_loc, self.n_kv_loc = h.n_heads // N, h.n_kv_heads // N
        self.q_loc, self.kv_loc, self.hid_loc = h.dim // N, h.kv_dim // N, h.hidden_dim // N
        self.vocab_loc = h.vocab_size // N if h.vocab_size % N == 0 else None

        def up(a: np.ndarray) -> torch.Tensor:
            return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(self.device)

        def rows(name, l, r0, n):
            return up(self.mf.f32(name, l)[r0:r0 + n])

        def cols(name, l, c0, n):
            return up(self.mf.f32(name, l)[:, c0:c0 + n])

        r = rank
        self.embedding = up(self.mf.f32("embedding"))
        self.layers = []
        for l in range(h.n_layers):
            self.layers.append({
                "qkv": torch.cat([rows("wq", l, r * self.q_loc, self.q_loc), rows("wk", l, r * self.kv_loc, self.kv_loc),
                                  rows("wv", l, r * self.kv_loc, self.kv_loc)], dim=0),
                "wo": cols("wo", l, r * self.q_loc, self.q_loc),
                "w1": rows("w1", l, r * self.hid_loc, self.hid_loc),
                "w3": rows("w3", l, r * self.hid_loc, self.hid_loc),
                "w2": cols("w2", l, r * self.hid_loc, self.hid_loc),
                "rms_att": up(self.mf.f32("rms_att", l)).reshape(-1),
                "rms_ffn": up(self.mf.f32("rms_ffn", l)).reshape(-1),
            })
        self.rms_final = up(self.mf.f32("rms_final")).reshape(-1)
        if self.vocab_loc is not None:
            self.wcls = rows("wcls", -1, r * self.vocab_loc, self.vocab_loc)
        else:                                                     # vocab not divisible: replicate the classifier
            self.wcls = up(self.mf.f32("wcls"))
        self.freqs = rope_frequencies(h).to(self.device)
        L = h.n_layers
        self.act_gelu = h.hidden_act != HIDDEN_ACT_SILU
        self.native = self.device.type == "cuda"
        if self.native:
            from ..ops import kernels as _K          # fails loudly if the CUDA library is missing on a GPU box
            _K.cuda_lib()
            self._K = _K
            self._init_native()
        else:
            self.k_cache = torch.zeros(n_slots, L, self.seq_len, self.n_kv_loc, hs, device=self.device)
            self.v_cache = torch.zeros(n_slots, L, self.seq_len, self.n_kv_loc, hs, device=self.device)
            self.pages = _Slots(n_slots)
            self.logits = torch.zeros(max_rows, h.vocab_size, device=self.device)

    # ---- native GPU path ---------------------------------------------------------------------------
    def _init_native(self):
        from ..parallel.symm import SymmetricBuffer
        from .llama import PAGE, PageAllocator
        K, h, dev, N, R = self._K, self.h, self.device, self.world, self.max_rows
        if h.dim % 32 or self.q_loc % 4 or self.hid_loc % 4 or (self.vocab_loc or h.vocab_size) % 4:
            raise ValueError("f32 GPU path: dim must be a multiple of 32, the per-rank widths multiples of 4")
        if self.vocab_loc is None and N > 1:
            raise ValueError("f32 GPU path with tensor parallelism needs vocab_size divisible by the number of ranks")
        self.max_pages = (self.seq_len + PAGE - 1) // PAGE
        self.n_pages = self.n_slots * self.max_pages
        self.pages = PageAllocator(self.n_pages, self.n_slots, self.max_pages)
        self.block_table = torch.zeros(self.n_slots, self.max_pages, dtype=torch.int32, device=dev)
        kv_shape = (self.n_pages, self.n_kv_loc, PAGE, self.hs)
        self.k_cache = [torch.zeros(kv_shape, dtype=torch.float32, device=dev) for _ in range(h.n_layers)]
        self.v_cache = [torch.zeros(kv_shape, dtype=torch.float32, device=dev) for _ in range(h.n_layers)]
        self.desc = torch.zeros(3, R, dtype=torch.int32, device=dev)          # tokens | pos | slot
        self.h_desc = torch.zeros(3, R, dtype=torch.int32).pin_memory()
        self.x = [torch.zeros(R, h.dim, device=dev) for _ in range(2)]
        self.y = torch.zeros(R, h.dim, device=dev)
        self.qkv = torch.zeros(R, self.q_loc + 2 * self.kv_loc, device=dev)
        self.q = torch.zeros(R, self.q_loc, device=dev)
        self.att = torch.zeros(R, self.q_loc, device=dev)
        self.g = torch.zeros(R, self.hid_loc, device=dev)
        self.u = torch.zeros(R, self.hid_loc, device=dev)
        self.hbuf = torch.zeros(R, self.hid_loc, device=dev)
        self.part = torch.zeros(R, h.dim, device=dev)
        self.lg_loc = torch.zeros(R, self.vocab_loc or h.vocab_size, device=dev)
        self.out_tokens = torch.zeros(R, dtype=torch.int32, device=dev)
        self.max_splits = 8
        self.att_scratch = torch.zeros(K.attention_scratch_bytes(R, self.n_heads_loc, self.hs, self.max_splits),
                                       dtype=torch.uint8, device=dev)
        # symmetric exchange area: [att partials N x R x dim][ffn partials][logits R x vocab][flags]
        part_bytes = N * R * h.dim * 4
        self._off_att, self._off_ffn, self._off_logits = 0, part_bytes, 2 * part_bytes
        self._off_flags = (2 * part_bytes + R * h.vocab_size * 4 + 255) // 256 * 256
        self.symm = SymmetricBuffer(self._off_flags + 4096, self.rank, N, self.group)
        sp, r = self.symm.ptrs, self.rank
        fl = [p + self._off_flags for p in sp]
        self._part_att = [p + self._off_att for p in sp]
        self._part_ffn = [p + self._off_ffn for p in sp]
        self._logit_ptrs = [p + self._off_logits for p in sp]
        self._flags_att_local, self._flags_ffn_local, self._flags_log_local = fl[r], fl[r] + 32, fl[r] + 64
        self._flags_att_peer = [p + 4 * r for p in fl]
        self._flags_ffn_peer = [p + 32 + 4 * r for p in fl]
        self._flags_log_peer = [p + 64 + 4 * r for p in fl]
        self.step_counter = fl[r] + 96
        self._counters = [fl[r] + 128 + 4 * i for i in range(3)]
        self.logits = self.symm.tensor(self._off_logits, (R, h.vocab_size), torch.float32)
        torch.cuda.synchronize()

    def _forward_native(self, tokens, positions, slots, logits_from: int, greedy: bool):
        """The launch sequence of one forward pass (reference graph: src/llm.cpp:185-432 without the Q80 casts)."""
        from .llama import _IntPtr, _Off
        K, h, N, L = self._K, self.h, self.world, self.h.n_layers
        M = len(tokens)
        hd = self.h_desc.numpy()
        hd[0, :M], hd[1, :M], hd[2, :M] = tokens, positions, slots
        for s_, p_ in zip(slots, positions):
            self.pages.ensure(int(s_), int(p_) + 1)
        if self.pages.dirty:
            self.block_table.copy_(torch.from_numpy(self.pages.table), non_blocking=False)
            self.pages.dirty = False
        self.desc.copy_(self.h_desc, non_blocking=True)
        tok, pos, slot = self.desc[0], self.desc[1], self.desc[2]
        sc, R, dim = self.step_counter, self.max_rows, h.dim
        stride = R * dim
        multi = N > 1
        K.embedding(self.embedding, tok, self.x[0], M, dim, _IntPtr(sc))
        cur = 0
        n_splits = max(1, min(self.max_splits, self.max_pages // 8))

        def norm(pending, flag_add, w, rows=M, x_off=0):
            """y = rmsnorm(x + merge of the pending partial sums) * w; the merged residual goes to the other x buffer."""
            nonlocal cur
            x_in = _Off(self.x[cur], x_off)
            if pending is None:
                K.reduce_norm_quant(x_in, None, None, 0, stride, None, sc, L, flag_add, w, h.norm_epsilon, dim, rows, rows,
                                    None, None, y_out=self.y)
                return
            parts, flags = pending
            K.reduce_norm_quant(x_in, _Off(self.x[cur ^ 1], x_off), parts + x_off, N, stride, flags if multi else None, sc, L,
                                flag_add, w, h.norm_epsilon, dim, rows, rows, None, None, y_out=self.y)
            cur ^= 1

        def push(partial, bufs, flags_peer, flag_add, counter):
            K.f32_push(partial, M, dim, bufs, self.rank * stride, dim, flags_peer if multi else None, sc, L, flag_add, counter)

        pending = None
        for l, W in enumerate(self.layers):
            norm(pending, l, W["rms_att"])
            K.gemm_f32(W["qkv"], self.y[:M], self.qkv[:M])
            K.f32_rope_kv(self.qkv, M, self.q_loc, self.kv_loc, self.hs, self.n_kv_loc, self.freqs, pos, slot, self.block_table,
                          self.max_pages, self.q, self.k_cache[l], self.v_cache[l])
            K.attention(self.q, M, self.n_heads_loc, self.n_kv_loc, self.hs, self.k_cache[l], self.v_cache[l], pos, slot,
                        self.block_table, self.max_pages, n_splits, self.att_scratch, None, None, M, y_out=self.att)
            K.gemm_f32(W["wo"], self.att[:M], self.part[:M])
            push(self.part, self._part_att, self._flags_att_peer, l + 1, self._counters[0])
            pending = (self._part_att[self.rank], self._flags_att_local)
            norm(pending, l + 1, W["rms_ffn"])
            K.gemm_f32(W["w1"], self.y[:M], self.g[:M])
            K.gemm_f32(W["w3"], self.y[:M], self.u[:M])
            K.f32_silu_mul(self.g[:M], self.u[:M], self.hbuf[:M], gelu=self.act_gelu)
            K.gemm_f32(W["w2"], self.hbuf[:M], self.part[:M])
            push(self.part, self._part_ffn, self._flags_ffn_peer, l + 1, self._counters[1])
            pending = (self._part_ffn[self.rank], self._flags_ffn_local)
        Ml = M - logits_from
        norm(pending, L, self.rms_final, rows=Ml, x_off=logits_from * dim * 4)
        vl = self.vocab_loc or h.vocab_size
        K.gemm_f32(self.wcls, self.y[:Ml], self.lg_loc[:Ml])
        K.f32_push(self.lg_loc, Ml, vl, self._logit_ptrs, self.rank * vl, h.vocab_size,
                   self._flags_log_peer if multi else None, sc, 1, 0, self._counters[2])
        if multi:
            K.wait_flags(self._flags_log_local, N, sc, 1, 0)
        if greedy:
            K.argmax(self.logits, Ml, h.vocab_size, h.vocab_size, self.out_tokens)
        return self.logits[:Ml]

    # ---- LlamaEngine surface ---------------------------------------------------------------------
    def alloc_slot(self) -> int:
        return self.pages.alloc_slot()

    def free_slot(self, slot: int):
        self.pages.free_slot(slot)

    def ensure_capacity(self, slot: int, n_tokens: int):
        if n_tokens > self.seq_len:
            raise RuntimeError(f"sequence of {n_tokens} tokens exceeds max_seq_len")
        if self.native:
            self.pages.ensure(slot, n_tokens)

    def weight_bytes(self) -> int:
        n = sum(t.numel() for L in self.layers for t in L.values()) + self.wcls.numel() + self.rms_final.numel()
        return 4 * n

    # ---- CPU fallback (no GPU on the machine): PyTorch ops, gloo all-reduce ----------------------------------------
    def _all_reduce(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t, group=self.group)
        return t

    def _mm(self, x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
        """x[M][n] @ w[d][n]^T in f32."""
        return x @ w.t()

    def _rms(self, x, w):
        return w * (x * torch.rsqrt((x * x).mean(-1, keepdim=True) + self.h.norm_epsilon))

    def _rope(self, x, pos):
        M = x.shape[0]
        xr = x.reshape(M, -1, self.hs // 2, 2)
        ang = pos.float()[:, None] * self.freqs[None, :]
        c, s = torch.cos(ang)[:, None, :], torch.sin(ang)[:, None, :]
        x0, x1 = xr[..., 0], xr[..., 1]
        return torch.stack([x0 * c - x1 * s, x0 * s + x1 * c], dim=-1).reshape(x.shape)

    @torch.no_grad()
    def forward(self, tokens: Sequence[int], positions: Sequence[int], slots: Sequence[int], logits_from: int = 0,
                greedy: bool = False) -> torch.Tensor:
        """Rows may mix sequences; rows of one sequence must be in increasing position order."""
        h, M = self.h, len(tokens)
        assert 1 <= M <= self.max_rows
        if self.native:
            return self._forward_native(tokens, positions, slots, logits_from, greedy)
        dev = self.device
        pos = torch.tensor(list(positions), device=dev, dtype=torch.long)
        slot_t = torch.tensor(list(slots), device=dev, dtype=torch.long)
        x = self.embedding[torch.tensor(list(tokens), device=dev, dtype=torch.long)]
        G = self.n_heads_loc // self.n_kv_loc
        # chunks = maximal runs of consecutive positions of one slot (a prompt chunk or one decode row)
        chunks, i = [], 0
        while i < M:
            j = i + 1
            while j < M and slots[j] == slots[i] and positions[j] == positions[j - 1] + 1:
                j += 1
            chunks.append((i, j))
            i = j
        scale = 1.0 / math.sqrt(self.hs)
        for l, W in enumerate(self.layers):
            y = self._rms(x, W["rms_att"])
            qkv = self._mm(y, W["qkv"])
            q = self._rope(qkv[:, :self.q_loc], pos).reshape(M, self.n_heads_loc, self.hs)
            k = self._rope(qkv[:, self.q_loc:self.q_loc + self.kv_loc], pos).reshape(M, self.n_kv_loc, self.hs)
            v = qkv[:, self.q_loc + self.kv_loc:].reshape(M, self.n_kv_loc, self.hs)
            self.k_cache[slot_t, l, pos] = k
            self.v_cache[slot_t, l, pos] = v
            att = torch.empty(M, self.n_heads_loc, self.hs, device=dev)
            for (a, b) in chunks:
                s, p0, p1 = int(slots[a]), int(positions[a]), int(positions[b - 1]) + 1
                kk = self.k_cache[s, l, :p1].repeat_interleave(G, dim=1)              # [S, heads, hs]
                vv = self.v_cache[s, l, :p1].repeat_interleave(G, dim=1)
                sc = torch.einsum("thd,shd->hts", q[a:b], kk) * scale
                mask = torch.arange(p1, device=dev)[None, :] > pos[a:b, None]         # causal inside the chunk
                sc = sc.masked_fill(mask[None], float("-inf"))
                att[a:b] = torch.einsum("hts,shd->thd", torch.softmax(sc, dim=-1), vv)
            x = x + self._all_reduce(self._mm(att.reshape(M, self.q_loc), W["wo"]))
            y = self._rms(x, W["rms_ffn"])
            g = self._mm(y, W["w1"])
            u = self._mm(y, W["w3"])
            act = torch.nn.functional.gelu(g, approximate="tanh") if self.act_gelu else torch.nn.functional.silu(g)
            x = x + self._all_reduce(self._mm(act * u, W["w2"]))
        y = self._rms(x[logits_from:], self.rms_final)
        lg = self._mm(y, self.wcls)
        if self.world > 1 and self.vocab_loc is not None:
            parts = [torch.empty_like(lg) for _ in range(self.world)]
            dist.all_gather(parts, lg, group=self.group)
            lg = torch.cat(parts, dim=-1)
        n_out = M - logits_from
        self.logits[:n_out] = lg
        return self.logits[:n_out]

    def prefill(self, slot: int, tokens: Sequence[int], start_pos: int = 0, chunk: Optional[int] = None):
        chunk = chunk or self.max_rows
        out = None
        for i in range(0, len(tokens), chunk):
            part = list(tokens[i:i + chunk])
            pos = list(range(start_pos + i, start_pos + i + len(part)))
            out = self.forward(part, pos, [slot] * len(part), logits_from=len(part) - 1)
        return out

    def decode_step_greedy(self, tokens, positions, slots, logits_from: int = 0) -> List[int]:
        if self.native:                                   # device argmax: only token ids cross PCIe
            self.forward(tokens, positions, slots, logits_from, greedy=True)
            return self.out_tokens[:len(tokens) - logits_from].cpu().tolist()
        lg = self.forward(tokens, positions, slots, logits_from)
        return [int(t) for t in torch.argmax(lg, dim=-1).cpu()]

    def decode_greedy_host(self, tokens, positions, slots) -> List[int]:
        return self.decode_step_greedy(tokens, positions, slots, 0)

    def close(self):
        if self.native:
            torch.cuda.synchronize()
            self.symm.close()
        self.mf.close()
