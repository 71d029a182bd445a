"""Thin torch-tensor wrappers over the C ABI (one function per kernel launch).

Every wrapper launches on ``torch.cuda.current_stream()`` so the calls compose with CUDA graph
capture; nothing here allocates except where stated.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import lib
from .lib import DlEpilogue, DlNormProlog, EPI_QKV_ROPE, EPI_SILU_MUL, EPI_STORE, check, cuda_lib

Q40T_TILE_BYTES = 9216
XQ_ROW_BYTES = 160
PAGE_SIZE = 32


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


# ---- buffer-size helpers ------------------------------------------------------------------------
def q40t_bytes(d: int, n: int) -> int:
    return int(cuda_lib().dl_q40t_bytes(d, n))


def xq_bytes(K: int, mcap: int) -> int:
    return (K // 128) * mcap * XQ_ROW_BYTES


def xh_bytes(K: int, ncap: int) -> int:
    assert ncap % 8 == 0
    return (K // 128) * (ncap // 8) * 2048


def alloc_xq(K: int, mcap: int, device) -> torch.Tensor:
    return torch.zeros(xq_bytes(K, mcap), dtype=torch.uint8, device=device)


def alloc_xh(K: int, ncap: int, device) -> torch.Tensor:
    return torch.zeros(xh_bytes(K, ncap), dtype=torch.uint8, device=device)


# ---- weights ----------------------------------------------------------------------------------------
def q40_repack(raw: torch.Tensor, rows: int, n: int, q40t: torch.Tensor, d_pad: int, dst_row0: int = 0,
               dst_row_step: int = 1):
    check(cuda_lib().dl_q40_repack(raw.data_ptr(), rows, n, q40t.data_ptr(), d_pad, dst_row0, dst_row_step,
                                   _stream()), "dl_q40_repack")
    lib.count_launch()


def q40t_dequant(q40t: torch.Tensor, d: int, n: int) -> torch.Tensor:
    out = torch.empty(d, n, dtype=torch.float32, device=q40t.device)
    check(cuda_lib().dl_q40t_dequant(q40t.data_ptr(), d, n, out.data_ptr(), _stream()), "dl_q40t_dequant")
    lib.count_launch()
    return out


def make_q40t(raw_q40: torch.Tensor, d: int, n: int) -> torch.Tensor:
    """raw Q40 file bytes of a [d][n] matrix (uint8 cuda tensor) -> Q40T image (allocates)."""
    d_pad = (d + 127) // 128 * 128
    out = torch.zeros(q40t_bytes(d, n), dtype=torch.uint8, device=raw_q40.device)
    q40_repack(raw_q40, d, n, out, d_pad)
    return out


# ---- epilogue builders ---------------------------------------------------------------------------
def epi_store(dst: Sequence[int], ld: int, dst_off: int = 0, q80_roundtrip: bool = False,
              flags: Optional[Sequence[int]] = None, step_counter: Optional[int] = None, flag_mul: int = 0,
              flag_add: int = 0, cta_counter: Optional[int] = None, residual: Optional[int] = None,
              residual_ld: int = 0, ll: bool = False, wire_q80: bool = False) -> DlEpilogue:
    """ll=True: dst[] hold LL words {f32 bits, epoch} (tensor-parallel partial sums); step_counter / flag_mul /
    flag_add define the epoch and no flags are released.  wire_q80=True (with flags): the partial sums travel as Q80
    rows (ld int8 + ld/32 fp16 scales) instead of f32; dst_off / ld keep their f32-layout meaning (row = offset / ld)."""
    e = DlEpilogue()
    e.kind = EPI_STORE
    assert not (ll and wire_q80)
    e.ll = 1 if ll else (2 if wire_q80 else 0)
    if wire_q80:
        assert q80_roundtrip and residual is None and dst_off % ld == 0
    if ll:
        assert flags is None and step_counter is not None
        e.step_counter = step_counter
        e.flag_mul = flag_mul
        e.flag_add = flag_add
    e.residual = residual
    e.residual_ld = residual_ld
    e.n_dst = len(dst)
    for i, p in enumerate(dst):
        e.dst[i] = p
    e.ld = ld
    e.dst_off = dst_off
    e.q80_roundtrip = 1 if q80_roundtrip else 0
    if flags is not None:
        assert len(flags) == len(dst) and step_counter is not None and cta_counter is not None
        for i, p in enumerate(flags):
            e.flag[i] = p
        e.step_counter = step_counter
        e.flag_mul = flag_mul
        e.flag_add = flag_add
        e.cta_counter = cta_counter
    return e


def epi_qkv_rope(q_out: torch.Tensor, q_dim: int, kv_dim: int, head_size: int, rope_freqs: torch.Tensor,
                 pos: torch.Tensor, slot: torch.Tensor, block_table: torch.Tensor, max_pages: int,
                 n_kv_heads: int, k_cache: torch.Tensor, v_cache: torch.Tensor) -> DlEpilogue:
    e = DlEpilogue()
    e.kind = EPI_QKV_ROPE
    e.q_out = q_out.data_ptr()
    e.q_dim, e.kv_dim, e.head_size = q_dim, kv_dim, head_size
    e.rope_freqs = rope_freqs.data_ptr()
    e.pos, e.slot, e.block_table = pos.data_ptr(), slot.data_ptr(), block_table.data_ptr()
    e.max_pages, e.page_size, e.n_kv_heads = max_pages, PAGE_SIZE, n_kv_heads
    e.k_cache, e.v_cache = k_cache.data_ptr(), v_cache.data_ptr()
    e.kv_f32 = 1 if k_cache.dtype == torch.float32 else 0
    return e


def epi_silu_mul(xq_out: Optional[torch.Tensor], xh_out: Optional[torch.Tensor], out_mcap: int,
                 gelu: bool = False) -> DlEpilogue:
    e = DlEpilogue()
    e.kind = EPI_SILU_MUL
    e.xq_out = _p(xq_out)
    e.xh_out = _p(xh_out)
    e.out_mcap = out_mcap
    e.act_gelu = 1 if gelu else 0
    return e


# ---- launches ------------------------------------------------------------------------------------
def reduce_norm_quant(x_in, x_out, partials, n_src, src_stride, flags, step_counter, flag_mul, flag_add, norm_w,
                      eps, dim, M, mcap, xq_out, xh_out, y_out=None, pdl=False, ll=False, wire_q80=False):
    mode = (1 if ll else (2 if wire_q80 else 0)) if (partials and n_src) else 0
    check(cuda_lib().dl_reduce_norm_quant(_p(x_in), _p(x_out), partials, n_src, src_stride, flags, step_counter,
                                          flag_mul, flag_add, mode, _p(norm_w), eps,
                                          dim, M, mcap, _p(xq_out),
                                          _p(xh_out), _p(y_out), 1 if pdl else 0, _stream()),
          "dl_reduce_norm_quant")
    lib.count_launch()


def quant_q80(x: torch.Tensor, M: int, K: int, mcap: int, xq_out=None, xh_out=None):
    check(cuda_lib().dl_quant_q80(x.data_ptr(), M, K, mcap, _p(xq_out), _p(xh_out), _stream()), "dl_quant_q80")
    lib.count_launch()


def embedding(table: torch.Tensor, tokens: torch.Tensor, x: torch.Tensor, M: int, dim: int, step_counter=None):
    check(cuda_lib().dl_embedding(table.data_ptr(), tokens.data_ptr(), x.data_ptr(), M, dim, _p(step_counter),
                                  _stream()), "dl_embedding")
    lib.count_launch()


def norm_prolog(x_in, x_out, partials: Optional[int], n_src: int, src_stride: int, flags: Optional[int],
                step_counter: Optional[int], flag_mul: int, flag_add: int, norm_w: torch.Tensor,
                eps: float, ll: bool = False) -> DlNormProlog:
    """Fused merge-add + RMSNorm + Q80 prologue of the GEMV (x_in/x_out: tensors or objects with data_ptr)."""
    p = DlNormProlog()
    p.x_in = _p(x_in)
    p.x_out = _p(x_out)
    p.partials = partials
    p.n_src = n_src if partials else 0
    p.src_stride = src_stride
    p.flags = flags
    p.step_counter = step_counter
    p.flag_mul, p.flag_add = flag_mul, flag_add
    p.norm_w = norm_w.data_ptr()
    p.eps = eps
    p.ll = 1 if (ll and partials) else 0
    return p


def gemv_q40(w: torch.Tensor, d: int, n: int, xq: Optional[torch.Tensor], M: int, epi: DlEpilogue,
             row_split: int = 1, k_split: int = 1, norm: Optional[DlNormProlog] = None, pdl: bool = False,
             tiles_per_cta: int = 1):
    """M<=8 rows.  Activations: XQ slabs with row capacity == M (xq), or the fused norm prologue."""
    check(cuda_lib().dl_gemv_q40(w.data_ptr(), d, n, _p(xq), M, row_split, k_split, tiles_per_cta,
                                 C.byref(norm) if norm is not None else None, C.byref(epi),
                                 1 if pdl else 0, _stream()), "dl_gemv_q40")
    lib.count_launch()


def gemm_q40_tc(w: torch.Tensor, d: int, n: int, xh: torch.Tensor, ncap: int, M: int, epi: DlEpilogue,
                k_split: int = 1, pdl: bool = False):
    check(cuda_lib().dl_gemm_q40_tc(w.data_ptr(), d, n, xh.data_ptr(), ncap, M, k_split, C.byref(epi),
                                    1 if pdl else 0, _stream()), "dl_gemm_q40_tc")
    lib.count_launch()


def wh_bytes(d: int, n: int) -> int:
    return int(cuda_lib().dl_wh_bytes(d, n))


def q40t_to_wh(q40t: torch.Tensor, d: int, n: int, out: torch.Tensor = None) -> torch.Tensor:
    """Q40T image -> WH (fp16 tiles in the UMMA core-matrix image) for gemm_wh_tc; `out` = caller's scratch."""
    if out is None:
        out = torch.empty(wh_bytes(d, n), dtype=torch.uint8, device=q40t.device)
    assert out.numel() >= wh_bytes(d, n)
    check(cuda_lib().dl_q40t_to_wh(q40t.data_ptr(), d, n, out.data_ptr(), _stream()), "dl_q40t_to_wh")
    lib.count_launch()
    return out


def gemm_wh_tc(wh: torch.Tensor, d: int, n: int, xh: torch.Tensor, ncap: int, M: int, epi: DlEpilogue,
               force_cg: int = 0, pdl: bool = False):
    """Prompt-sized batches: persistent fp16 SS-mode tcgen05 GEMM on pre-expanded weights (CTA pairs)."""
    check(cuda_lib().dl_gemm_wh_tc(wh.data_ptr(), d, n, xh.data_ptr(), ncap, M, C.byref(epi), force_cg,
                                   1 if pdl else 0, _stream()), "dl_gemm_wh_tc")
    lib.count_launch()


def attention_scratch_bytes(M: int, n_heads: int, head_size: int, max_splits: int) -> int:
    return int(cuda_lib().dl_attention_scratch_bytes(M, n_heads, head_size, max_splits))


def attention(q, M, n_heads, n_kv_heads, head_size, k_cache, v_cache, pos, slot, block_table, max_pages,
              n_splits, scratch, xq_out, xh_out, out_mcap, y_out=None, pdl=False, row_skip=None):
    check(cuda_lib().dl_attention(q.data_ptr(), M, n_heads, n_kv_heads, head_size, k_cache.data_ptr(),
                                  v_cache.data_ptr(), 1 if k_cache.dtype == torch.float32 else 0,
                                  pos.data_ptr(), slot.data_ptr(), block_table.data_ptr(), max_pages, PAGE_SIZE,
                                  n_splits, _p(scratch), _p(xq_out), _p(xh_out), out_mcap, _p(y_out), _p(row_skip),
                                  1 if pdl else 0, _stream()), "dl_attention")
    lib.count_launch()


def attention_prefill(q, M, n_heads, n_kv_heads, head_size, k_cache, v_cache, pos, slot, block_table, max_pages,
                      meta, xq_out, xh_out, out_mcap, y_out=None, pdl=False, tc=False):
    """Tensor-core causal attention of prompt chunks: meta = [n_seg, begin0, end0, ...] (device int32), each
    range being consecutive positions of one sequence.  tc=True: the tcgen05 / TMEM kernel (head size 128)."""
    assert k_cache.dtype == torch.float16
    args = (q.data_ptr(), M, n_heads, n_kv_heads, head_size, k_cache.data_ptr(), v_cache.data_ptr(), pos.data_ptr(),
            slot.data_ptr(), block_table.data_ptr(), max_pages, PAGE_SIZE, meta.data_ptr(), _p(xq_out), _p(xh_out), out_mcap,
            _p(y_out))
    if tc:
        check(cuda_lib().dl_attention_prefill_tc(*args, int(k_cache.shape[0]), 1 if pdl else 0, _stream()), "dl_attention_prefill_tc")
    else:
        check(cuda_lib().dl_attention_prefill(*args, 1 if pdl else 0, _stream()), "dl_attention_prefill")
    lib.count_launch()


def argmax(logits, M: int, vocab: int, ld: int, out: torch.Tensor, flags=None, n_flags=0, step_counter=None,
           flag_mul=0, flag_add=0, pos_inc=None, history=None, history_cap=0, scratch=None, pdl=False):
    lp = logits.data_ptr() if isinstance(logits, torch.Tensor) else logits
    if scratch is None:
        scratch = _argmax_scratch(M, out.device)
    check(cuda_lib().dl_argmax(lp, M, vocab, ld, out.data_ptr(), flags, n_flags, step_counter, flag_mul, flag_add,
                               _p(pos_inc), _p(history), history_cap, scratch.data_ptr(), 1 if pdl else 0,
                               _stream()), "dl_argmax")
    lib.count_launch()


_ARGMAX_SCRATCH = {}


def _argmax_scratch(M: int, device) -> torch.Tensor:
    """Shared default scratch (zeroed once; the kernel leaves its counters at zero)."""
    key = str(device)
    t = _ARGMAX_SCRATCH.get(key)
    need = int(cuda_lib().dl_argmax_scratch_bytes(256))
    if t is None:
        t = torch.zeros(need, dtype=torch.uint8, device=device)
        _ARGMAX_SCRATCH[key] = t
    return t


def wait_flags(flags: int, n: int, step_counter: int, flag_mul: int, flag_add: int):
    check(cuda_lib().dl_wait_flags(flags, n, step_counter, flag_mul, flag_add, _stream()), "dl_wait_flags")
    lib.count_launch()


def sample_rows(logits: torch.Tensor, temperature: torch.Tensor, topp: torch.Tensor, coin: torch.Tensor,
                tokens_out: torch.Tensor):
    """tokens_out[m] = sample(logits[m]) on the device; temperature / topp / coin are f32 device tensors [M]."""
    M, vocab = logits.shape
    check(cuda_lib().dl_sample_rows(logits.data_ptr(), M, vocab, logits.stride(0), temperature.data_ptr(), topp.data_ptr(),
                                    coin.data_ptr(), tokens_out.data_ptr(), _stream()), "dl_sample_rows")
    lib.count_launch()


def gemm_f32(w: torch.Tensor, x: torch.Tensor, out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    """out[M][d] (+)= x[M][n] @ w[d][n]^T with plain f32 FMAs (the f32-weights path).  w may be a column
    slice of a wider matrix (row stride w.stride(0)); x and out are row-major."""
    d, n = w.shape
    M = x.shape[0]
    assert w.dtype == torch.float32 and x.dtype == torch.float32 and w.stride(1) == 1 and x.stride(1) == 1
    if out is None:
        out = torch.empty(M, d, device=x.device, dtype=torch.float32)
        accumulate = False
    check(cuda_lib().dl_gemm_f32(w.data_ptr(), d, n, w.stride(0), x.data_ptr(), M, x.stride(0), out.data_ptr(),
                                 out.stride(0), 1 if accumulate else 0, _stream()), "dl_gemm_f32")
    lib.count_launch((M + 7) // 8)
    return out


def f32_rope_kv(qkv: torch.Tensor, M: int, q_dim: int, kv_dim: int, head_size: int, n_kv_heads: int, freqs, pos, slot,
                block_table, max_pages: int, q_out: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor):
    """f32 path: RoPE on q / k, q -> q_out, k / v -> paged KV cache."""
    check(cuda_lib().dl_f32_rope_kv(qkv.data_ptr(), M, q_dim, kv_dim, head_size, n_kv_heads, freqs.data_ptr(), pos.data_ptr(),
                                    slot.data_ptr(), block_table.data_ptr(), max_pages, q_out.data_ptr(), k_cache.data_ptr(),
                                    v_cache.data_ptr(), 1 if k_cache.dtype == torch.float32 else 0, _stream()), "dl_f32_rope_kv")
    lib.count_launch()


def f32_silu_mul(g: torch.Tensor, u: torch.Tensor, out: torch.Tensor, gelu: bool = False):
    check(cuda_lib().dl_f32_silu_mul(g.data_ptr(), u.data_ptr(), out.data_ptr(), g.numel(), 1 if gelu else 0, _stream()),
          "dl_f32_silu_mul")
    lib.count_launch()


def f32_push(src: torch.Tensor, M: int, w: int, dst: Sequence[int], dst_off: int, ld_dst: int,
             flags: Optional[Sequence[int]] = None, step_counter: Optional[int] = None, flag_mul: int = 0, flag_add: int = 0,
             cta_counter: Optional[int] = None):
    """f32 path: this rank's slice -> every rank's exchange buffer (+ release flags)."""
    n = len(dst)
    d = (C.c_void_p * n)(*dst)
    f = (C.c_void_p * n)(*flags) if flags is not None else None
    check(cuda_lib().dl_f32_push(src.data_ptr(), M, w, src.stride(0), d, n, dst_off, ld_dst, f, step_counter, flag_mul, flag_add,
                                 cta_counter, _stream()), "dl_f32_push")
    lib.count_launch()


MEGA_MAX_ROWS = 4
MEGA_LL_WORDS = 9          # 8-byte LL words per 32-value Q80 block (8 x {4 int8, epoch} + {fp16 scale, epoch})


def mega_ll_bytes(world: int, dim: int) -> int:
    """Bytes of one exchange buffer of the persistent decode kernel: [world][4 rows][dim/32][9] LL words."""
    return world * MEGA_MAX_ROWS * (dim // 32) * MEGA_LL_WORDS * 8


def decode_mega_supported(args: "lib.DlDecodeArgs", grid: int) -> bool:
    return cuda_lib().dl_decode_mega_supported(C.byref(args), grid) == 0


def decode_mega(args: "lib.DlDecodeArgs", grid: int):
    """One persistent cooperative launch = one whole forward pass for args.M <= 4 rows."""
    check(cuda_lib().dl_decode_mega(C.byref(args), grid, _stream()), "dl_decode_mega")
    lib.count_launch()


def decode_mega_plan(d: int, n: int, align: int, grid: int, cta: int):
    """(first row, end row, pipeline chunks) of CTA `cta` for a [d][n] matmul phase (host helper)."""
    r0, r1, nc = C.c_int(), C.c_int(), C.c_int()
    check(cuda_lib().dl_decode_mega_plan(d, n, align, grid, cta, C.byref(r0), C.byref(r1), C.byref(nc)),
          "dl_decode_mega_plan")
    return r0.value, r1.value, nc.value


def choose_tc_split(d: int, n: int, M: int, n_sms: int) -> int:
    """Cluster split-K factor of the tcgen05 GEMM: a power of two (<= 8) that fills the chip without
    ever launching more CTAs than SMs (one CTA per SM: a second wave of half-empty CTAs costs more
    than the extra parallelism buys).  The activation tile is 16/32/64/128 rows (256 with a 256-row
    capacity), so wide batches already bring ceil(M / tile) CTAs per row tile."""
    row_tiles = (d + 127) // 128
    nt = 16 if M <= 16 else 32 if M <= 32 else 64 if M <= 64 else 128
    m_tiles = (M + nt - 1) // nt
    # wide tiles reduce one 32-row chunk at a time over DSMEM (two cluster syncs per chunk): only worth it when
    # every CTA still keeps a long k-loop (measured: w2 with 112 k-tiles gains 1.5x at S=4, qkv/wo with 32 lose)
    min_kt = 1 if M <= 32 else 24
    S = 1
    while row_tiles * m_tiles * S * 2 <= n_sms and S < 8 and (n // 128) // (S * 2) >= min_kt:
        S *= 2
    return S


def choose_splits(d: int, n: int, epi_kind: int, n_sms: int = 148, M: int = 1):
    """Work split of the GEMV -> (row_split R, k_split S, tiles_per_cta T); grid = ceil(units/T)*S.

    The kernel runs one CTA per SM, so the grid is kept within one wave:
      * more row tiles than SMs: every CTA streams T = ceil(row_tiles/SMs) tiles back to back;
      * fewer: the K range of each tile is cut over a cluster of S CTAs (DSMEM reduction), and only
        if that still leaves most SMs idle are tiles also cut into R row parts (which shrinks the
        pipeline stages; R<=2 when the epilogue needs 64-row groups)."""
    row_tiles = (d + 127) // 128
    kt = n // 128
    if row_tiles >= n_sms:
        return 1, 1, (row_tiles + n_sms - 1) // n_sms
    # clusters must be co-scheduled inside one GPC (16-20 SMs): power-of-two sizes pack best, and the
    # grid is kept at <= ~128 CTAs so that placement never spills into a second wave
    budget = (n_sms * 7) // 8
    S = 1
    while S * 2 <= min(8, kt) and row_tiles * S * 2 <= budget:
        S *= 2
    while S < min(8, kt) and int(cuda_lib().dl_gemv_q40_smem_bytes(n, M, S)) > 200 * 1024:
        S *= 2
    R = 1
    max_r = 2 if epi_kind == EPI_SILU_MUL else 4
    while R < max_r and row_tiles * S * R * 2 <= budget:
        R *= 2
    return R, S, 1
