"""GPU inference engine: Llama forward on sm_100a kernels, tensor-parallel over peer memory.

This replaces the reference's graph builder + executor + device (buildLlmNet, src/llm.cpp:126-438;
NnExecutor::forward, src/nn/nn-executor.cpp:57-192; NnCpuDevice, src/nn/nn-cpu.cpp) with a fixed
sequence of five fused kernels per layer, captured once per batch shape into a CUDA graph:

    K1 reduce_norm_quant   merge-add of the peers' partial sums + residual + RMSNorm + Q80 cast
    K2 qkv matmul          merged q|k|v weight; epilogue: RoPE + paged KV-cache write
    K3 attention           paged GQA flash-decode; epilogue: Q80 cast
    K4 wo matmul           epilogue: Q80-cast partial sum pushed to every peer GPU + release flags
    K5 (K1 again for ffn)  ... w1|w3 matmul (epilogue SiLU*mul + Q80) ... w2 matmul (push)

Tensor parallelism follows the reference's slicing (src/llm.cpp:131-142): q/k/v/w1/w3/wcls are
split along the output dim, wo/w2 along the input dim, KV cache by kv head; the residual stream is
replicated.  The socket all-gather + merge-add (SYNC_NODE_SLICES + OP_MERGE_ADD) becomes peer
stores from the matmul epilogue plus a flag wait in the next kernel's prologue.
"""
from __future__ import annotations

import os

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..formats.mfile import ModelFile, HIDDEN_ACT_SILU
from ..ops import kernels as K
from ..ops import quant
from ..ops.lib import EPI_QKV_ROPE, EPI_SILU_MUL, EPI_STORE
from ..parallel.symm import SymmetricBuffer, validate_tp
from .oracle import rope_frequencies

GEMV_MAX_M = 8              # rows the dp4a GEMV kernel supports (XQ buffer capacity)
# rows up to which the GEMV path is USED: its dp4a work grows with every row (w13: 19 us at M=1, 33 us at M=4)
# while the tcgen05 GEMM costs the same for 1..16 rows (33 us).  Measured decode step of the 8B model: 3 users
# 2.54 ms (GEMV) vs 2.59 ms (tcgen05), 6 users 4.03 vs 2.72 ms, 8 users 4.68 vs 2.75 ms -> cross over above 3 rows
GEMV_PATH_MAX_M = int(os.environ.get("DLLAMA_GEMV_MAX_M", "3"))
LL_MAX_ROWS = 32             # batches up to this many rows use the LL exchange (see LlamaEngine.__init__)
LL_MAX_WORLD = 4             # largest tensor-parallel group that exchanges partial sums as LL words
PREFILL_MAX_SEG = 8          # prompt-chunk segments per batch (attention_prefill.cu: PF_MAX_SEG)
# Prompt-sized batches run on pre-expanded fp16 weights ("WH", csrc/cuda/gemm_wh_tc.cu): a plain SS-mode tcgen05 GEMM with
# CTA pairs instead of the fused dequantising one.  Measured on the 8B shapes (TFLOP/s, fused kernel -> WH kernel):
# w13 M=512 756 -> 897, M=1024 845 -> 1079, M=2048 908 -> 1193; qkv M=1024 623 -> 730, M=2048 832 -> 1061;
# wo / w2 break even at 1024 rows and gain 17-21 % at 2048.  The expansion itself runs at 4.3-6.0 TB/s (w13: 50 us).
WH_MIN_ROWS = int(os.environ.get("DLLAMA_WH_MIN_ROWS", "768"))             # resident copy: free to use
WH_MIN_ROWS_SCRATCH = int(os.environ.get("DLLAMA_WH_MIN_ROWS_SCRATCH", "2048"))   # per-call expansion: ~13 % of the GEMM time at 2048 rows
# batches from this many rows run their prompt chunks on the tcgen05 attention kernel (attention_prefill_tc.cu, 128-row
# tiles, head size 128); smaller ones on the mma.sync kernel with 64-row tiles
PREFILL_TC_MIN_ROWS = int(os.environ.get("DLLAMA_PREFILL_TC_MIN_ROWS", "128"))
WIRE_Q80 = os.environ.get("DLLAMA_WIRE_Q80", "1") != "0"
PAGE = K.PAGE_SIZE


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class PageAllocator:
    """Host-side KV page allocator: per-slot page lists over one shared pool (per-user KV caches;
    the reference has a single cache shared by all users, src/llm.cpp:197-198)."""

    def __init__(self, n_pages: int, n_slots: int, max_pages_per_seq: int):
        self.free_pages = list(range(n_pages - 1, -1, -1))
        self.free_slots = list(range(n_slots - 1, -1, -1))
        self.max_pages = max_pages_per_seq
        self.table = np.zeros((n_slots, max_pages_per_seq), dtype=np.int32)
        self.n_used = np.zeros(n_slots, dtype=np.int32)
        self.dirty = True

    def alloc_slot(self) -> int:
        if not self.free_slots:
            raise RuntimeError("no free sequence slot")
        return self.free_slots.pop()

    def free_slot(self, slot: int):
        for i in range(int(self.n_used[slot])):
            self.free_pages.append(int(self.table[slot, i]))
        self.n_used[slot] = 0
        self.free_slots.append(slot)

    def release_pages(self, slot: int):
        """Return a slot's pages to the pool; the slot id itself stays with its owner (the scheduler)."""
        for i in range(int(self.n_used[slot])):
            self.free_pages.append(int(self.table[slot, i]))
        self.n_used[slot] = 0

    def ensure(self, slot: int, n_tokens: int):
        need = (n_tokens + PAGE - 1) // PAGE
        if need > self.max_pages:
            raise RuntimeError(f"sequence of {n_tokens} tokens exceeds max_seq_len")
        while self.n_used[slot] < need:
            if not self.free_pages:
                raise RuntimeError("KV cache exhausted")
            self.table[slot, self.n_used[slot]] = self.free_pages.pop()
            self.n_used[slot] += 1
            self.dirty = True


class LlamaEngine:
    def __init__(self, model_path: str, rank: int = 0, world: int = 1, group=None, max_rows: int = 32,
                 n_slots: int = 4, max_seq_len: int = 2048, kv_dtype: torch.dtype = torch.float16,
                 sync_q80: bool = True, use_pdl: bool = True, use_graphs: bool = True, n_pages: Optional[int] = None,
                 device: Optional[torch.device] = None, verbose: bool = False, bucket_rows: bool = False,
                 use_mega: Optional[bool] = None, prefill_wh: str = "auto"):
        self.mf = ModelFile(model_path, max_seq_len=max_seq_len)
        h = self.h = self.mf.header
        if h.weight_type != quant.Q40:
            raise NotImplementedError("GPU engine: only Q40 weights (+Q80 activations) are wired up; "
                                      "f32 models run through models.oracle")
        validate_tp(world, h.n_kv_heads, h.dim, h.hidden_dim, h.n_heads, h.vocab_size)
        self.rank, self.world, self.group = rank, world, group
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.max_rows = max_rows
        self.use_pdl, self.use_graphs = use_pdl, use_graphs
        self.sync_q80 = sync_q80
        # Decode batches of <= 4 rows can run as ONE persistent cooperative kernel (csrc/cuda/decode_mega.cu) instead of the
        # launch sequence below.  Measured on 8B single-stream decode (tok/s, launch sequence vs persistent kernel):
        # 1 GPU 593 / 477, 2 GPUs 628 / 573, 4 GPUs 575 / 555, 8 GPUs 464 / 573 -- a step is a chain of ~8 grid-wide
        # hand-offs per layer either way; the persistent kernel wins where the launch sequence's fence + counter + flag
        # exchange dominates.  Default: on for world >= 8.  DLLAMA_MEGA=1 / 0 or use_mega=True / False force it.
        env_mega = os.environ.get("DLLAMA_MEGA", "auto")
        if use_mega is None:
            use_mega = (world >= 8) if env_mega == "auto" else (env_mega != "0")
        self.use_mega = bool(use_mega)
        self.seq_len = min(h.seq_len, max_seq_len)
        self.n_sms = K.cuda_lib().dl_device_sm_count()
        dev = self.device

        # ---- per-rank geometry (reference: sliceRowMatmul / sliceColMatmul / sliceKvCache) -------
        N = world
        self.dim, self.hs = h.dim, h.head_size
        self.n_heads_loc = h.n_heads // N
        self.n_kv_loc = h.n_kv_heads // N
        self.q_loc = self.n_heads_loc * self.hs
        self.kv_loc = self.n_kv_loc * self.hs
        self.hid_loc = h.hidden_dim // N
        self.vocab_loc = h.vocab_size // N
        self.qkv_d = self.q_loc + 2 * self.kv_loc

        # ---- weights ---------------------------------------------------------------------------
        self._load_weights(verbose)
        self._init_wh(os.environ.get("DLLAMA_PREFILL_WH", prefill_wh))

        # ---- activations -----------------------------------------------------------------------
        R = max_rows
        self.ncap = max(16, _round_up(R, 128) if R > 64 else (16 if R <= 16 else 32 if R <= 32 else 64))
        # batch descriptor: tokens | pos | slot | row_skip | meta.  Runs of >= 16 consecutive positions of one
        # sequence (prompt chunks) are listed in meta = [n_seg, begin0, end0, ...] and served by the tensor-core
        # prefill attention; row_skip marks their rows for the per-row kernel, which serves everything else.
        DW = max(R, 2 * PREFILL_MAX_SEG + 1)
        self.desc = torch.zeros(5, DW, dtype=torch.int32, device=dev)
        self.h_desc = torch.zeros(5, DW, dtype=torch.int32).pin_memory()
        self.tokens, self.pos, self.slot, self.row_skip, self.meta = (self.desc[i] for i in range(5))
        self._hd = self.h_desc.numpy()                                   # host view, filled by set_batch
        self._desc3, self._h_desc3 = self.desc[:3], self.h_desc[:3]      # tokens | pos | slot: all a batch without prompt chunks needs
        self.prefill_attention = kv_dtype == torch.float16 and self.hs in (64, 128)
        self._chunk = False
        self.x = [torch.zeros(R, h.dim, dtype=torch.float32, device=dev) for _ in range(2)]
        self.q = torch.zeros(R, self.q_loc, dtype=torch.float32, device=dev)
        self.xq_dim = K.alloc_xq(h.dim, GEMV_MAX_M, dev)
        self.xq_att = K.alloc_xq(self.q_loc, GEMV_MAX_M, dev)
        self.xq_hid = K.alloc_xq(self.hid_loc, GEMV_MAX_M, dev)
        self.xh_dim = K.alloc_xh(h.dim, self.ncap, dev)
        self.xh_att = K.alloc_xh(self.q_loc, self.ncap, dev)
        self.xh_hid = K.alloc_xh(self.hid_loc, self.ncap, dev)
        self.rope_freqs = rope_frequencies(h).to(dev)
        self.out_tokens = torch.zeros(R, dtype=torch.int32, device=dev)
        self.h_out_tokens = torch.zeros(R, dtype=torch.int32).pin_memory()
        self.history_cap = 4096
        self.history = torch.zeros(self.history_cap * R, dtype=torch.int32, device=dev)

        # ---- symmetric exchange area: partial sums, logits, flags, counters ------------------------
        # Tensor-parallel partial sums travel as LL words {f32 bits, epoch} (8 bytes per value, one atomic store):
        # the consumer polls the data itself, so the hand-off needs no fence, counter or flag (DLLAMA_LL=0 keeps
        # the plain f32 + release-flag protocol for comparison)
        # Measured (8B decode): TP=2 518 -> 634 tok/s with LL; at TP=8 the 8 x 2x larger reads and the polling of
        # every consumer CTA cost more than the fence they save (545 -> 460), so wide groups keep the flag protocol.
        # DLLAMA_LL=1 / 0 forces it on / off.
        ll_env = os.environ.get("DLLAMA_LL", "auto")
        self.use_ll = N > 1 and (ll_env == "1" or (ll_env != "0" and N <= LL_MAX_WORLD))
        # LL words for latency-bound batches (<= LL_MAX_ROWS rows: decode), plain f32 + flags for prompt-sized
        # batches (there the 2x bytes cost more than the fence: TTFT 4.8 -> 5.5 ms at TP=4); the two protocols
        # use disjoint regions so stale data of one can never be read as the other
        ll_rows = min(R, LL_MAX_ROWS) if self.use_ll else 0
        part_bytes = N * R * h.dim * 4
        ll_bytes = N * ll_rows * h.dim * 8
        logit_bytes = R * h.vocab_size * 4
        self._off_att, self._off_ffn = 0, part_bytes
        self._off_att_ll, self._off_ffn_ll = 2 * part_bytes, 2 * part_bytes + ll_bytes
        self._off_logits = 2 * part_bytes + 2 * ll_bytes
        self._off_flags = _round_up(self._off_logits + logit_bytes, 256)
        # persistent decode kernel: Q80 LL-word exchange buffers (wo / w2 partial sums, greedy candidates)
        mega_ll = K.mega_ll_bytes(N, h.dim)
        self._off_mega_att = self._off_flags + 4096
        self._off_mega_ffn = self._off_mega_att + mega_ll
        self._off_mega_tok = self._off_mega_ffn + mega_ll
        total = self._off_mega_tok + _round_up(N * K.MEGA_MAX_ROWS * 2 * 8, 256)
        self.symm = SymmetricBuffer(total, rank, world, group)
        sp = self.symm.ptrs
        self.part_att = [p + self._off_att for p in sp]
        self.part_ffn = [p + self._off_ffn for p in sp]
        self.part_att_ll = [p + self._off_att_ll for p in sp]
        self.part_ffn_ll = [p + self._off_ffn_ll for p in sp]
        self._ll_rows = ll_rows
        self.logits_ptrs = [p + self._off_logits for p in sp]
        # flag words: [att: 8][ffn: 8][logits: 8][step counter][cta counters: 8]
        fl = [p + self._off_flags for p in sp]
        self.flags_att_local, self.flags_ffn_local, self.flags_log_local = fl[rank], fl[rank] + 32, fl[rank] + 64
        self.flags_att_peer = [p + 4 * rank for p in fl]
        self.flags_ffn_peer = [p + 32 + 4 * rank for p in fl]
        self.flags_log_peer = [p + 64 + 4 * rank for p in fl]
        self.step_counter = fl[rank] + 96
        self.cta_counters = [fl[rank] + 128 + 4 * i for i in range(8)]
        self.logits = self.symm.tensor(self._off_logits, (R, h.vocab_size), torch.float32)

        # ---- KV cache ----------------------------------------------------------------------------
        self.max_pages = (self.seq_len + PAGE - 1) // PAGE
        # bucket_rows (serving): every batch is padded to a power-of-two row count with rows of a hidden
        # scratch sequence, so continuous batching replays a handful of CUDA graphs instead of capturing
        # a new one for every (rows, first-logit-row) combination it produces
        self.bucket_rows = bucket_rows
        self.n_slots = n_slots
        n_slots = n_slots + (1 if bucket_rows else 0)
        self.n_pages = (n_pages or self.n_slots * self.max_pages) + (1 if bucket_rows else 0)
        # the page pool is shared by all sequences and grows on demand, so it does not have to cover
        # n_slots x max_seq_len (Llama-3.1 headers say 131072): cap it at 85 % of the free HBM
        page_bytes = 2 * h.n_layers * self.n_kv_loc * PAGE * self.hs * torch.empty((), dtype=kv_dtype).element_size()
        if n_pages is None and dev.type == "cuda":
            free_b, _ = torch.cuda.mem_get_info(dev)
            fit = int(0.85 * free_b) // page_bytes
            if fit < self.n_pages:
                if fit < self.max_pages + 1:
                    raise RuntimeError(f"KV cache: not even one full sequence of {self.seq_len} tokens fits in "
                                       f"{free_b >> 20} MiB of free device memory; lower --max-seq-len")
                print(f"💡 KV page pool capped at {fit} pages ({fit * page_bytes >> 20} MiB) instead of "
                      f"{self.n_pages} (n_slots x max_seq_len); sequences share it on demand", flush=True)
                self.n_pages = fit
        if world > 1 and n_pages is None and dev.type == "cuda":
            # every rank replays the same batches with the same page tables: agree on the smallest pool
            import torch.distributed as dist
            t = torch.tensor([self.n_pages], dtype=torch.int64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
            self.n_pages = int(t.item())
        self.pages = PageAllocator(self.n_pages, n_slots, self.max_pages)
        self.block_table = torch.zeros(n_slots, self.max_pages, dtype=torch.int32, device=dev)
        self.h_block_table = torch.zeros(n_slots, self.max_pages, dtype=torch.int32).pin_memory()
        kv_shape = (self.n_pages, self.n_kv_loc, PAGE, self.hs)
        self.k_cache = [torch.zeros(kv_shape, dtype=kv_dtype, device=dev) for _ in range(h.n_layers)]
        self.v_cache = [torch.zeros(kv_shape, dtype=kv_dtype, device=dev) for _ in range(h.n_layers)]
        self._pad_slot, self._Mp = -1, 0
        if bucket_rows:
            self._pad_slot = n_slots - 1
            self.pages.free_slots.remove(self._pad_slot)
            self.pages.ensure(self._pad_slot, 1)
        self.max_splits = 32
        # split-KV partials exist only for batches small enough to leave SMs idle (M * splits(M) <= ~n_sms / n_kv):
        # size the scratch for the largest M * splits product, not for max_rows x max_splits (2 GB at 4096 rows)
        att_rows = max(m * self._attention_splits(m) for m in range(1, min(R, self.n_sms) + 1))
        self.att_scratch = torch.zeros(K.attention_scratch_bytes(max(att_rows, R), self.n_heads_loc, self.hs, 1),
                                       dtype=torch.uint8, device=dev)
        self.argmax_scratch = torch.zeros(int(K.cuda_lib().dl_argmax_scratch_bytes(256)), dtype=torch.uint8,
                                          device=dev)
        self._init_mega(kv_dtype)
        self._staging_ev: Optional[torch.cuda.Event] = None
        self._graphs: Dict[Tuple, torch.cuda.CUDAGraph] = {}
        self._graph_launches: Dict[Tuple, int] = {}
        self._graph_stream = torch.cuda.Stream(device=dev)
        torch.cuda.synchronize()

    # ------------------------------------------------------------------------------------------------
    # persistent decode kernel (csrc/cuda/decode_mega.cu)
    # ------------------------------------------------------------------------------------------------
    def _init_mega(self, kv_dtype):
        h, dev, N = self.h, self.device, self.world
        self._mega_ok_cache: Dict[int, bool] = {}
        self.mega_grid = 0
        if not self.use_mega or dev.type != "cuda" or kv_dtype != torch.float16 or not self.sync_q80:
            self.use_mega = False
            return
        from ..ops.lib import DlDecodeLayer
        R4 = K.MEGA_MAX_ROWS
        self.mega_grid = min(self.n_sms, int(os.environ.get("DLLAMA_MEGA_CTAS", str(self.n_sms))))
        tab = (DlDecodeLayer * h.n_layers)()
        for l, W in enumerate(self.layers):
            tab[l].qkv, tab[l].wo, tab[l].w13, tab[l].w2 = (W[k].data_ptr() for k in ("qkv", "wo", "w13", "w2"))
            tab[l].rms_att, tab[l].rms_ffn = W["rms_att"].data_ptr(), W["rms_ffn"].data_ptr()
            tab[l].k_cache, tab[l].v_cache = self.k_cache[l].data_ptr(), self.v_cache[l].data_ptr()
        raw = np.frombuffer(bytes(tab), dtype=np.uint8).copy()
        self._mega_layers = torch.from_numpy(raw).to(dev)
        self._mega_x = torch.zeros(2, R4, h.dim, dtype=torch.float32, device=dev)
        self._mega_q = torch.zeros(R4, self.q_loc, dtype=torch.float32, device=dev)
        self._mega_h = torch.zeros(R4, self.hid_loc, dtype=torch.float32, device=dev)
        self._mega_att_o = torch.zeros(R4 * self.n_heads_loc * self.max_splits * self.hs, dtype=torch.float32, device=dev)
        self._mega_att_ml = torch.zeros(R4 * self.n_heads_loc * self.max_splits * 2, dtype=torch.float32, device=dev)
        self._mega_best = torch.zeros(self.mega_grid * R4 * 2, dtype=torch.float32, device=dev)
        self._mega_bar = torch.zeros(64, dtype=torch.int32, device=dev)     # barrier words [0, 2), merger flags [16, 32)
        self._mega_out = torch.zeros(R4, dtype=torch.int32, device=dev)
        # a time-out inside the kernel traps (the context is lost); the reason is left in mapped host memory
        self._mega_debug = torch.zeros(16, dtype=torch.int32).pin_memory()
        self._mega_trace: Optional[torch.Tensor] = None
        self._mega_trace_cta = 0
        self._mega_pf_dist = int(os.environ.get("DLLAMA_MEGA_PF", "0"))
        self._mega_timeout_ns = int(float(os.environ.get("DLLAMA_MEGA_TIMEOUT_S", "20")) * 1e9)

    def _mega_args(self, M: int, row0: int, greedy: bool, device_loop: bool):
        from ..ops.lib import DlDecodeArgs
        h, N = self.h, self.world
        a = DlDecodeArgs()
        a.n_layers, a.dim, a.q_dim, a.kv_dim, a.hid = h.n_layers, h.dim, self.q_loc, self.kv_loc, self.hid_loc
        a.vocab_loc, a.vocab, a.head_size = self.vocab_loc, h.vocab_size, self.hs
        a.n_heads, a.n_kv_heads = self.n_heads_loc, self.n_kv_loc
        a.rank, a.world = self.rank, N
        a.eps = h.norm_epsilon
        a.act_gelu = 0 if h.hidden_act == HIDDEN_ACT_SILU else 1
        a.q80_sync = 1 if self.sync_q80 else 0
        a.layers = self._mega_layers.data_ptr()
        a.rms_final, a.wcls = self.rms_final.data_ptr(), self.wcls.data_ptr()
        a.embedding, a.rope_freqs = self.embedding.data_ptr(), self.rope_freqs.data_ptr()
        a.M, a.row0 = M, row0
        a.tokens, a.pos, a.slot = self.tokens.data_ptr(), self.pos.data_ptr(), self.slot.data_ptr()
        a.block_table, a.max_pages = self.block_table.data_ptr(), self.max_pages
        a.x[0], a.x[1] = self._mega_x[0].data_ptr(), self._mega_x[1].data_ptr()
        a.q, a.h = self._mega_q.data_ptr(), self._mega_h.data_ptr()
        a.att_o, a.att_ml = self._mega_att_o.data_ptr(), self._mega_att_ml.data_ptr()
        a.max_splits = self.max_splits
        a.att_splits = max(1, min(self.max_splits, self.mega_grid // max(1, M * self.n_kv_loc)))
        a.cta_best = self._mega_best.data_ptr()
        a.grid_bar = self._mega_bar.data_ptr()
        a.step_counter = self.step_counter
        sp = self.symm.ptrs
        for p in range(N):
            a.ll_att[p] = sp[p] + self._off_mega_att
            a.ll_ffn[p] = sp[p] + self._off_mega_ffn
            a.ll_tok[p] = sp[p] + self._off_mega_tok
            a.logits[p] = self.logits_ptrs[p]
            a.flag_logits[p] = self.flags_log_peer[p]
        a.greedy = 1 if greedy else 0
        a.tokens_out = self.out_tokens.data_ptr()
        a.advance = 1 if device_loop else 0
        a.history = self.history.data_ptr() if device_loop else None
        a.history_cap = self.history_cap
        a.timeout_ns = self._mega_timeout_ns
        a.debug = self._mega_debug.data_ptr()
        a.trace = self._mega_trace.data_ptr() if self._mega_trace is not None else None
        a.trace_cta = self._mega_trace_cta
        a.pf_dist = self._mega_pf_dist
        a.dbg = int(os.environ.get("DLLAMA_MEGA_DBG", "0"))
        return a

    def _mega_usable(self, M: int, chunk: bool) -> bool:
        if not self.use_mega or chunk or M > K.MEGA_MAX_ROWS:
            return False
        ok = self._mega_ok_cache.get(M)
        if ok is None:
            ok = K.decode_mega_supported(self._mega_args(M, 0, True, False), self.mega_grid)
            self._mega_ok_cache[M] = ok
        return ok

    def enable_mega_trace(self, on: bool = True, cta: int = 0):
        """Per-phase globaltimer stamps of one CTA (16 per layer + 2): read with mega_trace_ns() after a forward.
        Graphs captured before the switch keep their old setting."""
        self._mega_trace = torch.zeros(self.h.n_layers * 32 + 32, dtype=torch.int64, device=self.device) if on else None
        self._mega_trace_cta = cta
        self._graphs.clear()

    def mega_trace_ns(self) -> np.ndarray:
        return self._mega_trace.cpu().numpy().astype(np.int64)

    # ------------------------------------------------------------------------------------------------
    def _load_weights(self, verbose: bool):
        h, mf, dev, r, N = self.h, self.mf, self.device, self.rank, self.world
        nb_dim = h.dim // 32

        def up(arr: np.ndarray) -> torch.Tensor:
            t = torch.empty(arr.size, dtype=torch.uint8)
            t.numpy()[:] = arr.reshape(-1)
            return t.to(dev, non_blocking=False)

        def rows(name, l, row0, nrows):
            return up(mf.raw_rows(name, l, row0, nrows))

        def cols(name, l, col0, ncols):
            e = mf.entry(name, l)
            nb = e.cols // 32
            v = mf.raw(name, l).reshape(e.rows, nb, 18)[:, col0 // 32:(col0 + ncols) // 32, :]
            return up(np.ascontiguousarray(v))

        def f32(name, l=-1):
            return torch.from_numpy(mf.f32(name, l).reshape(-1).copy()).to(dev)

        self.embedding = torch.empty(h.vocab_size, h.dim, dtype=torch.float32, device=dev)
        e = mf.entry("embedding")
        chunk = 16384
        for r0 in range(0, h.vocab_size, chunk):
            n = min(chunk, h.vocab_size - r0)
            a = mf.data[e.offset + r0 * h.dim * 4: e.offset + (r0 + n) * h.dim * 4].view(np.float32)
            self.embedding[r0:r0 + n] = torch.from_numpy(a.reshape(n, h.dim).copy()).to(dev)
        self.layers = []
        for l in range(h.n_layers):
            L = {}
            qkv = torch.zeros(K.q40t_bytes(self.qkv_d, h.dim), dtype=torch.uint8, device=dev)
            dpad = _round_up(self.qkv_d, 128)
            K.q40_repack(rows("wq", l, r * self.q_loc, self.q_loc), self.q_loc, h.dim, qkv, dpad, 0, 1)
            K.q40_repack(rows("wk", l, r * self.kv_loc, self.kv_loc), self.kv_loc, h.dim, qkv, dpad, self.q_loc, 1)
            K.q40_repack(rows("wv", l, r * self.kv_loc, self.kv_loc), self.kv_loc, h.dim, qkv, dpad,
                         self.q_loc + self.kv_loc, 1)
            L["qkv"] = qkv
            wo = torch.zeros(K.q40t_bytes(h.dim, self.q_loc), dtype=torch.uint8, device=dev)
            K.q40_repack(cols("wo", l, r * self.q_loc, self.q_loc), h.dim, self.q_loc, wo, _round_up(h.dim, 128))
            L["wo"] = wo
            w13 = torch.zeros(K.q40t_bytes(2 * self.hid_loc, h.dim), dtype=torch.uint8, device=dev)
            dpad13 = _round_up(2 * self.hid_loc, 128)
            K.q40_repack(rows("w1", l, r * self.hid_loc, self.hid_loc), self.hid_loc, h.dim, w13, dpad13, 0, 2)
            K.q40_repack(rows("w3", l, r * self.hid_loc, self.hid_loc), self.hid_loc, h.dim, w13, dpad13, 1, 2)
            L["w13"] = w13
            w2 = torch.zeros(K.q40t_bytes(h.dim, self.hid_loc), dtype=torch.uint8, device=dev)
            K.q40_repack(cols("w2", l, r * self.hid_loc, self.hid_loc), h.dim, self.hid_loc, w2, _round_up(h.dim, 128))
            L["w2"] = w2
            L["rms_att"], L["rms_ffn"] = f32("rms_att", l), f32("rms_ffn", l)
            self.layers.append(L)
            torch.cuda.synchronize()
            if verbose and rank_zero(self):
                print(f"🕒 layer {l + 1}/{h.n_layers} loaded", flush=True)
        self.rms_final = f32("rms_final")
        wcls = torch.zeros(K.q40t_bytes(self.vocab_loc, h.dim), dtype=torch.uint8, device=dev)
        K.q40_repack(rows("wcls", -1, r * self.vocab_loc, self.vocab_loc), self.vocab_loc, h.dim, wcls,
                     _round_up(self.vocab_loc, 128))
        self.wcls = wcls
        torch.cuda.synchronize()

    def _init_wh(self, mode: str):
        """fp16 copies of the layer matrices for prompt-sized batches.  'resident': expanded once and kept (8B: 14 GB of
        the 180 GB); 'scratch': one buffer of the largest matrix, expanded before every use; 'off'; 'auto': resident
        when the engine can see batches that large and the copies take less than a quarter of the free memory."""
        h = self.h
        shapes = {"qkv": (self.qkv_d, h.dim), "wo": (h.dim, self.q_loc), "w13": (2 * self.hid_loc, h.dim), "w2": (h.dim, self.hid_loc)}
        self._wh_shapes = shapes
        self._wh, self._wh_scratch, self.wh_mode = {}, None, "off"
        if mode == "off" or self.device.type != "cuda" or self.max_rows < min(WH_MIN_ROWS, WH_MIN_ROWS_SCRATCH):
            return
        per_layer = sum(K.wh_bytes(d, n) for d, n in shapes.values())
        if mode == "auto":
            free_b, _ = torch.cuda.mem_get_info(self.device)
            mode = "resident" if per_layer * h.n_layers <= free_b // 4 else "scratch"
        if mode == "resident" and self.max_rows >= WH_MIN_ROWS:
            for L in self.layers:
                for name, (d, n) in shapes.items():
                    self._wh[L[name].data_ptr()] = K.q40t_to_wh(L[name], d, n)
            torch.cuda.synchronize()
            self.wh_mode = "resident"
        elif mode == "scratch" and self.max_rows >= WH_MIN_ROWS_SCRATCH:
            self._wh_scratch = torch.empty(max(K.wh_bytes(d, n) for d, n in shapes.values()), dtype=torch.uint8, device=self.device)
            self.wh_mode = "scratch"

    def wh_bytes(self) -> int:
        return sum(t.numel() for t in self._wh.values()) + (self._wh_scratch.numel() if self._wh_scratch is not None else 0)

    def weight_bytes(self) -> int:
        """Q40T bytes streamed per forward on this rank (roofline numerator)."""
        n = sum(t.numel() for L in self.layers for k, t in L.items() if t.dtype == torch.uint8)
        return n + self.wcls.numel()

    # ------------------------------------------------------------------------------------------------
    # slots / pages
    # ------------------------------------------------------------------------------------------------
    def usable_pages(self) -> int:
        """Pages the scheduler may hand out to requests (the padding sequence of bucket_rows keeps one)."""
        return self.n_pages - (1 if self.bucket_rows else 0)

    def alloc_slot(self) -> int:
        return self.pages.alloc_slot()

    def free_slot(self, slot: int):
        self.pages.free_slot(slot)

    def ensure_capacity(self, slot: int, n_tokens: int):
        self.pages.ensure(slot, n_tokens)

    def _sync_block_table(self):
        if self.pages.dirty:
            self.h_block_table.numpy()[:] = self.pages.table
            self.block_table.copy_(self.h_block_table, non_blocking=True)
            self.pages.dirty = False

    # ------------------------------------------------------------------------------------------------
    # the forward pass (launch sequence)
    # ------------------------------------------------------------------------------------------------
    def _matmul(self, w, d, n, M, xq, xh, epi, kind, norm=None):
        if M <= GEMV_PATH_MAX_M:
            R, S, T = K.choose_splits(d, n, kind, self.n_sms, M)
            K.gemv_q40(w, d, n, None if norm is not None else xq, M, epi, R, S, norm=norm, pdl=self.use_pdl,
                       tiles_per_cta=T)
        elif self.wh_mode == "resident" and M >= WH_MIN_ROWS and w.data_ptr() in self._wh:
            K.gemm_wh_tc(self._wh[w.data_ptr()], d, n, xh, self.ncap, M, epi, pdl=self.use_pdl)
        elif self.wh_mode == "scratch" and M >= WH_MIN_ROWS_SCRATCH and w is not self.wcls:
            K.q40t_to_wh(w, d, n, self._wh_scratch)      # plain launch: starts after the previous user of the scratch has finished
            K.gemm_wh_tc(self._wh_scratch, d, n, xh, self.ncap, M, epi, pdl=self.use_pdl)
        else:
            assert norm is None
            S = K.choose_tc_split(d, n, M, self.n_sms)
            K.gemm_q40_tc(w, d, n, xh, self.ncap, M, epi, S, pdl=self.use_pdl)

    def _launch_forward(self, M: int, row0: int, greedy: bool, device_loop: bool, chunk: bool = False):
        """Enqueue one forward over rows 0..M-1 of the descriptor; logits for rows row0..M-1."""
        h, N, pdl = self.h, self.world, self.use_pdl
        if self._mega_usable(M, chunk):
            K.decode_mega(self._mega_args(M, row0, greedy, device_loop), self.mega_grid)
            if not greedy and N > 1:
                K.wait_flags(self.flags_log_local, N, self.step_counter, 1, 0)
            return
        gemv = M <= GEMV_PATH_MAX_M
        xq = lambda t: t if gemv else None
        xh = lambda t: None if gemv else t
        mcap = M if gemv else self.ncap               # XQ slabs are written with row capacity == M
        L = h.n_layers
        multi = N > 1
        ll_fwd = self.use_ll and M <= self._ll_rows          # this forward exchanges LL words (same choice on every rank)
        # batches on the tensor-core path that use the flag protocol ship Q80 rows (int8 + fp16 scales) instead of f32:
        # 3.8x fewer NVLink bytes, which is what bounds prompt-sized exchanges (DLLAMA_WIRE_Q80=0: f32 as before)
        wire = (N > 1) and not gemv and not ll_fwd and self.sync_q80 and WIRE_Q80
        src_stride = (self._ll_rows if ll_fwd else self.max_rows) * h.dim
        part_att = self.part_att_ll if ll_fwd else self.part_att
        part_ffn = self.part_ffn_ll if ll_fwd else self.part_ffn
        sc = self.step_counter
        eps = h.norm_epsilon
        # Residual bookkeeping.  TP>1 (and the tensor-core path): matmul epilogues push Q80-cast partial
        # sums to every rank's exchange buffer; the next norm merges them (x ping-pongs between two
        # buffers).  TP=1 GEMV path: the wo/w2 epilogues add the Q80-cast product into x in place.
        inplace = gemv and not multi
        K.embedding(self.embedding, self.tokens, self.x[0], M, h.dim, _IntPtr(sc))
        cur = 0
        n_splits = self._attention_splits(M)

        def norm_then(w_norm, pending, flag_add, weight, d, kind, epi, x_off=0, rows=M):
            """RMSNorm(+merge of `pending` partials) feeding a dim->d matmul."""
            nonlocal cur
            merge = pending is not None and not inplace
            x_in = _Off(self.x[cur], x_off)
            x_out = _Off(self.x[cur ^ 1], x_off) if merge else None
            ll = ll_fwd and merge
            wq = wire and merge
            if wq:       # Q80 wire rows: dim int8 + dim/32 fp16 scales per row
                parts = pending[0][self.rank] + (x_off // (h.dim * 4)) * (h.dim + h.dim // 16)
            else:
                parts = (pending[0][self.rank] + x_off * (2 if ll else 1)) if merge else None
            flags = pending[1] if (merge and multi and not ll) else None
            fuse = False
            if rows <= GEMV_PATH_MAX_M and not wq:
                fuse = True          # the grid never exceeds one CTA per SM: the repeated prologue is cheap
            if fuse:
                nm = K.norm_prolog(x_in, x_out, parts, N, src_stride, flags, sc, L, flag_add, w_norm, eps, ll=ll)
                self._matmul(weight, d, h.dim, rows, None, None, epi, kind, norm=nm)
            elif rows <= GEMV_PATH_MAX_M:
                K.reduce_norm_quant(x_in, x_out, parts, N if merge else 0, src_stride, flags, sc, L, flag_add, w_norm,
                                    eps, h.dim, rows, rows, self.xq_dim, None, pdl=pdl, ll=ll, wire_q80=wq)
                self._matmul(weight, d, h.dim, rows, self.xq_dim, None, epi, kind)
            else:
                K.reduce_norm_quant(x_in, x_out, parts, N if merge else 0, src_stride, flags, sc, L, flag_add, w_norm,
                                    eps, h.dim, rows, self.ncap, None, self.xh_dim, pdl=pdl, ll=ll, wire_q80=wq)
                self._matmul(weight, d, h.dim, rows, None, self.xh_dim, epi, kind)
            if merge:
                cur ^= 1

        def push(buf_ptrs, flags_peer, flag_add, counter):
            if inplace:
                xp = self.x[cur].data_ptr()
                return K.epi_store([xp], h.dim, 0, self.sync_q80, residual=xp, residual_ld=h.dim)
            if ll_fwd:
                return K.epi_store(buf_ptrs, h.dim, self.rank * src_stride, self.sync_q80, None, sc, L, flag_add,
                                   ll=True)
            return K.epi_store(buf_ptrs, h.dim, self.rank * src_stride, self.sync_q80,
                               flags_peer if multi else None, sc, L, flag_add, counter, wire_q80=wire)

        pending = None
        for l, W in enumerate(self.layers):
            # --- attention block ---
            epi = K.epi_qkv_rope(self.q, self.q_loc, self.kv_loc, self.hs, self.rope_freqs, self.pos, self.slot,
                                 self.block_table, self.max_pages, self.n_kv_loc, self.k_cache[l], self.v_cache[l])
            norm_then(W["rms_att"], pending, l, W["qkv"], self.qkv_d, EPI_QKV_ROPE, epi)
            if chunk != 2:
                K.attention(self.q, M, self.n_heads_loc, self.n_kv_loc, self.hs, self.k_cache[l], self.v_cache[l], self.pos,
                            self.slot, self.block_table, self.max_pages, n_splits, self.att_scratch, xq(self.xq_att),
                            xh(self.xh_att), mcap, pdl=pdl, row_skip=self.row_skip if chunk else None)
            if chunk:        # prompt-chunk rows (skipped above) on the tensor cores
                K.attention_prefill(self.q, M, self.n_heads_loc, self.n_kv_loc, self.hs, self.k_cache[l], self.v_cache[l],
                                    self.pos, self.slot, self.block_table, self.max_pages, self.meta, xq(self.xq_att),
                                    xh(self.xh_att), mcap, pdl=pdl, tc=self.hs == 128 and M >= PREFILL_TC_MIN_ROWS)
            self._matmul(W["wo"], h.dim, self.q_loc, M, self.xq_att, self.xh_att,
                         push(part_att, self.flags_att_peer, l + 1, self.cta_counters[0]), EPI_STORE)
            pending = (part_att, self.flags_att_local)
            # --- feed-forward block ---
            epi = K.epi_silu_mul(xq(self.xq_hid), xh(self.xh_hid), mcap, gelu=(h.hidden_act != HIDDEN_ACT_SILU))
            norm_then(W["rms_ffn"], pending, l + 1, W["w13"], 2 * self.hid_loc, EPI_SILU_MUL, epi)
            self._matmul(W["w2"], h.dim, self.hid_loc, M, self.xq_hid, self.xh_hid,
                         push(part_ffn, self.flags_ffn_peer, l + 1, self.cta_counters[1]), EPI_STORE)
            pending = (part_ffn, self.flags_ffn_local)
        # --- final norm + logits for rows row0.. ---
        Ml = M - row0
        epi = K.epi_store(self.logits_ptrs, h.vocab_size, self.rank * self.vocab_loc, False,
                          self.flags_log_peer if multi else None, sc, 1, 0, self.cta_counters[2])
        norm_then(self.rms_final, pending, L, self.wcls, self.vocab_loc, EPI_STORE, epi, x_off=row0 * h.dim * 4,
                  rows=Ml)
        if greedy:
            K.argmax(self.logits, Ml, h.vocab_size, h.vocab_size,
                     self.tokens if device_loop else self.out_tokens,
                     self.flags_log_local if multi else None, N, sc, 1, 0,
                     pos_inc=self.pos if device_loop else None,
                     history=self.history if device_loop else None, history_cap=self.history_cap,
                     scratch=self.argmax_scratch, pdl=pdl)
        elif multi:
            K.wait_flags(self.flags_log_local, N, sc, 1, 0)

    def _attention_splits(self, M: int) -> int:
        ctas = M * self.n_kv_loc
        want = max(1, self.n_sms // max(1, ctas))
        max_by_ctx = max(1, self.max_pages // 8)           # >= 8 pages per split; the kernel uses fewer splits for short rows
        return int(min(want, max_by_ctx, self.max_splits))

    def _run(self, M: int, row0: int, greedy: bool, device_loop: bool = False, chunk: Optional[bool] = None):
        chunk = (0 if device_loop else self._chunk) if chunk is None else chunk
        key = (M, row0, greedy, device_loop, chunk)
        if not self.use_graphs:
            self._launch_forward(M, row0, greedy, device_loop, chunk)
            return
        g = self._graphs.get(key)
        if g is None:
            # warm up once eagerly (sets func attributes, loads modules), then capture
            self._launch_forward(M, row0, greedy and not device_loop, False, chunk)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            n0 = K.lib.launch_count
            with torch.cuda.graph(g, stream=self._graph_stream):
                self._launch_forward(M, row0, greedy, device_loop, chunk)
            self._graph_launches[key] = K.lib.launch_count - n0
            K.lib.launch_count = n0          # capture enqueued nothing
            self._graphs[key] = g
        g.replay()
        K.lib.count_launch(self._graph_launches[key])

    def _run_batch(self, M: int, logits_from: int, greedy: bool) -> int:
        """Run the batch staged by set_batch(); returns the index of the first wanted row in
        self.logits / self.out_tokens.  With bucket_rows the graph key is (padded rows, 0 or last row)."""
        if not self.bucket_rows:
            self._run(M, logits_from, greedy)
            return 0
        Mp = self._Mp
        if logits_from == M - 1:              # prompt chunk or single decode row: only the last row's logits
            self._run(Mp, Mp - 1, greedy)
            return 0
        tail = self._tail_rows(Mp)
        if tail and M - logits_from <= tail:  # prompt rows + a few sampling rows: logits of the last `tail` rows only
            self._run(Mp, Mp - tail, greedy)
            return tail - (M - logits_from)
        self._run(Mp, 0, greedy)              # logits of every row; the caller's rows are the tail
        return (Mp - M) + logits_from

    def _bucket(self, M: int) -> int:
        """Padded row count of a bucket_rows batch: powers of two up to 256, then multiples of 256 (a 1091-row batch
        runs as 1280 rows, not 2048)."""
        if M <= 256:
            Mp = 1
            while Mp < M:
                Mp *= 2
        else:
            Mp = _round_up(M, 256)
        return min(Mp, self.max_rows)

    def _buckets(self) -> List[int]:
        out, m = [], 1
        while m < min(256, self.max_rows):
            out.append(m)
            m *= 2
        out += list(range(256, self.max_rows + 1, 256)) if self.max_rows >= 256 else []
        if not out or out[-1] != self.max_rows:
            out.append(self.max_rows)
        return out

    def _tail_rows(self, Mp: int) -> int:
        """Large mixed batches (prompt rows of many requests + their sampling rows): the wcls matmul, the logits write
        and the argmax run on the last `tail` rows (>= one row per sequence slot) instead of on every row."""
        tail = 1
        while tail < min(self.n_slots, 256):
            tail *= 2
        return tail if Mp >= 2 * tail else 0

    def warmup_graphs(self):
        """Capture every CUDA graph a bucket_rows engine can ask for (power-of-two row counts x {all-row
        logits, last-row logits} x {argmax, logits}) on rows of the scratch sequence, so that no request
        ever waits for a capture.  Every rank must call it at the same point (it runs real forwards)."""
        assert self.bucket_rows and self.use_graphs
        hd = self.h_desc.numpy()
        for Mp in self._buckets():
            hd[0, :Mp] = 0
            hd[1, :Mp] = 0
            hd[2, :Mp] = self._pad_slot
            self.desc.copy_(self.h_desc, non_blocking=True)
            tail = self._tail_rows(Mp)
            row0s = sorted({0, Mp - 1} | ({Mp - tail} if tail else set()))
            if Mp <= 256:                                    # decode-sized batches without prompt chunks
                for row0 in row0s:
                    for greedy in (False, True):
                        self._run(Mp, row0, greedy, chunk=False)
            if self.prefill_attention and Mp >= 16:          # batches that contain prompt chunks
                hd[3, :Mp] = 0
                hd[4, 0] = 0                                 # no segment: both kernels run, the prefill one exits at once
                self.desc.copy_(self.h_desc, non_blocking=True)
                for row0 in row0s:
                    for greedy in (False, True):
                        self._run(Mp, row0, greedy, chunk=True)
        torch.cuda.synchronize()

    # ------------------------------------------------------------------------------------------------
    # public API
    # ------------------------------------------------------------------------------------------------
    def set_batch(self, tokens: Sequence[int], positions: Sequence[int], slots: Sequence[int]):
        """Stage the batch descriptor: pinned host -> device (the reference's control packet +
        token/position pipes, src/app.cpp:169-209)."""
        M = len(tokens)
        assert 1 <= M <= self.max_rows
        hd = self._hd
        # The pinned staging buffers (descriptor, block table) are read by the copy engine when the async copy
        # EXECUTES, which can be a whole forward pass after it was enqueued (prompt chunks and worker ranks run
        # back to back without a host sync).  Do not overwrite them before the previous copy has been consumed.
        ev = self._staging_ev
        if ev is not None:
            ev.synchronize()
        if M == 1 and not self.bucket_rows:          # single-stream decode: the latency-critical host path
            hd[0, 0], hd[1, 0], hd[2, 0] = tokens[0], positions[0], slots[0]
            self._chunk = False
            self.pages.ensure(int(slots[0]), int(positions[0]) + 1)
            if self.pages.dirty:
                self._sync_block_table()
            self._desc3.copy_(self._h_desc3, non_blocking=True)
            self._mark_staging()
            return 1
        off = 0
        if self.bucket_rows:
            Mp = self._bucket(M)
            off = Mp - M                      # padding rows first, real rows last (the last row stays last)
            hd[0, :off] = 0
            hd[1, :off] = 0
            hd[2, :off] = self._pad_slot
            self._Mp = Mp
        hd[0, off:off + M] = tokens
        hd[1, off:off + M] = positions
        hd[2, off:off + M] = slots
        # prompt chunks: maximal runs of consecutive positions of one sequence, >= 16 rows
        segs = []
        if self.prefill_attention and M >= 16:
            i = 0
            while i < M:
                j = i + 1
                while j < M and slots[j] == slots[i] and positions[j] == positions[j - 1] + 1:
                    j += 1
                if j - i >= 16 and len(segs) < PREFILL_MAX_SEG:
                    segs.append((off + i, off + j))
                i = j
        # 0 = no prompt chunk, 1 = chunks mixed with other rows, 2 = every row belongs to a chunk (a long prompt run as
        # one batch): then the per-row kernel has nothing to do and is not launched
        self._chunk = (2 if (off == 0 and sum(e - b for b, e in segs) == M) else 1) if segs else 0
        if segs:
            hd[3, :off + M] = 0
            hd[4, 0] = len(segs)
            for k, (b, e) in enumerate(segs):
                hd[3, b:e] = 1
                hd[4, 1 + 2 * k] = b
                hd[4, 2 + 2 * k] = e
        need = {}
        for s, p in zip(slots, positions):           # one ensure per sequence (a 4096-row prompt is one sequence)
            if p >= need.get(s, -1):
                need[s] = p
        for s, p in need.items():
            self.pages.ensure(int(s), int(p) + 1)
        self._sync_block_table()
        if segs:
            self.desc.copy_(self.h_desc, non_blocking=True)
        else:
            self._desc3.copy_(self._h_desc3, non_blocking=True)
        self._mark_staging()
        return M

    def _mark_staging(self):
        """Record the point in the stream after which the pinned staging buffers may be rewritten."""
        if self.device.type != "cuda":
            return
        if self._staging_ev is None:
            self._staging_ev = torch.cuda.Event()
        self._staging_ev.record()

    def forward(self, tokens, positions, slots, logits_from: int = 0, greedy: bool = False) -> torch.Tensor:
        """One forward over a batch of rows (any mix of decode rows and prefill-chunk rows; rows of
        the same sequence must be in increasing position order).  Returns the device logits view
        [M - logits_from][vocab] (valid until the next forward)."""
        M = self.set_batch(tokens, positions, slots)
        o = self._run_batch(M, logits_from, greedy)
        return self.logits[o:o + M - logits_from]

    def prefill(self, slot: int, tokens: Sequence[int], start_pos: int = 0, chunk: Optional[int] = None):
        """Evaluate a prompt in chunks of <= max_rows rows; returns logits of the last token."""
        chunk = chunk or self.max_rows
        out = None
        for i in range(0, len(tokens), chunk):
            part = list(tokens[i:i + chunk])
            pos = list(range(start_pos + i, start_pos + i + len(part)))
            out = self.forward(part, pos, [slot] * len(part), logits_from=len(part) - 1)
        return out

    def decode_step_greedy(self, tokens, positions, slots, logits_from: int = 0) -> List[int]:
        """One greedy step end to end for rows logits_from..M-1: H2D descriptor, forward, device
        argmax, D2H token ids."""
        M = self.set_batch(tokens, positions, slots)
        o = self._run_batch(M, logits_from, True)
        self.h_out_tokens.copy_(self.out_tokens, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.h_out_tokens[o:o + M - logits_from].tolist()

    def decode_greedy_host(self, tokens: List[int], positions: List[int], slots: List[int]) -> List[int]:
        return self.decode_step_greedy(tokens, positions, slots, 0)

    def decode_greedy_device(self, n_steps: int, M: int):
        """Device-resident greedy loop: each step's argmax feeds the next step's embedding and
        advances the positions without a host round trip.  Requires set_batch() first and enough
        KV pages for n_steps more tokens (ensure_capacity)."""
        assert not self.bucket_rows, "the device-resident loop runs exact batch shapes"
        for _ in range(n_steps):
            self._run(M, 0, True, device_loop=True)

    def read_history(self, n_steps: int, M: int, first_step: int) -> np.ndarray:
        """Tokens produced by the last device-loop steps: [n_steps][M]."""
        raw = self.history.cpu().numpy()
        out = np.zeros((n_steps, M), dtype=np.int32)
        for i in range(n_steps):
            idx = (first_step + i) % self.history_cap
            out[i] = raw[idx * M: idx * M + M]
        return out

    def current_step(self) -> int:
        t = self.symm.tensor(self._off_flags + 96, (1,), torch.int32)
        return int(t.item())

    def close(self):
        self._graphs.clear()
        torch.cuda.synchronize()
        self.symm.close()
        self.mf.close()


def rank_zero(engine) -> bool:
    return engine.rank == 0


class _IntPtr:
    """Lets a raw device address travel through wrappers that call .data_ptr()."""

    def __init__(self, ptr: int):
        self._ptr = ptr

    def data_ptr(self) -> int:
        return self._ptr


def _Off(t: torch.Tensor, byte_off: int) -> _IntPtr:
    return _IntPtr(t.data_ptr() + byte_off)
