"""Numerics of every CUDA kernel against plain PyTorch/NumPy f32 references of the same op.

Tolerances follow the reference's own tests (src/nn/nn-cpu-ops-test.cpp:240,289;
src/nn/nn-vulkan-test.cpp:581) but are much tighter where the arithmetic is exact.
"""
import math

import numpy as np
import pytest
import torch

from dllama_b200.ops import quant

pytestmark = pytest.mark.gpu


def K():
    from dllama_b200.ops import kernels
    return kernels


def _rand_q40(d, n, seed=0, scale=0.05):
    rng = np.random.default_rng(seed)
    w = (rng.standard_normal((d, n), dtype=np.float32) * scale).astype(np.float32)
    raw = quant.quantize_q40(w)                       # [d*n/32, 18]
    wdq = quant.dequantize_q40(raw).reshape(d, n)
    return raw.reshape(-1), wdq


def _decode_xq(xq_bytes, K_, mcap, M):
    a = xq_bytes.cpu().numpy().reshape(K_ // 128, mcap, 160)
    q = a[:, :M, :128].copy().view(np.int8).astype(np.float32)                  # [kt, M, 128]
    d = a[:, :M, 128:144].copy().view(np.float32)                               # [kt, M, 4]
    s = a[:, :M, 144:160].copy().view(np.int32)
    x = (q.reshape(K_ // 128, M, 4, 32) * d[..., None]).transpose(1, 0, 2, 3).reshape(M, K_)
    return x, q, d, s


def _decode_xh(xh_bytes, K_, ncap, M):
    """XH slabs -> f32 [M][K] in natural element order."""
    a = xh_bytes.cpu().numpy().view(np.float16).reshape(K_ // 128, ncap // 8, 16, 8, 8).astype(np.float32)
    out = np.zeros((M, K_), dtype=np.float32)
    perm = [0, 2, 1, 3]
    for e in range(32):
        hi, e2 = e >> 4, e & 15
        w, r = e2 >> 2, e2 & 3
        s = hi * 4 + perm[r]
        for kb in range(4):
            # [kt, group, chunk, row, slot]
            v = a[:, :, kb * 4 + w, :, s]                         # [kt, groups, 8]
            v = v.reshape(K_ // 128, ncap)[:, :M]                 # [kt, M]
            out[:, np.arange(K_ // 128) * 128 + kb * 32 + e] = v.T
    return out


def test_repack_dequant_roundtrip():
    k = K()
    for d, n in ((128, 128), (200, 256), (384, 1024)):
        raw, wdq = _rand_q40(d, n, seed=d)
        t = k.make_q40t(torch.from_numpy(raw).cuda(), d, n)
        out = k.q40t_dequant(t, d, n).cpu().numpy()
        np.testing.assert_array_equal(out, wdq)


def test_repack_interleave_and_offsets():
    k = K()
    d, n = 64, 256
    raw1, w1 = _rand_q40(d, n, 1)
    raw3, w3 = _rand_q40(d, n, 2)
    t = torch.zeros(k.q40t_bytes(2 * d, n), dtype=torch.uint8, device="cuda")
    k.q40_repack(torch.from_numpy(raw1).cuda(), d, n, t, 128, 0, 2)
    k.q40_repack(torch.from_numpy(raw3).cuda(), d, n, t, 128, 1, 2)
    out = k.q40t_dequant(t, 2 * d, n).cpu().numpy()
    np.testing.assert_array_equal(out[0::2], w1)
    np.testing.assert_array_equal(out[1::2], w3)


@pytest.mark.parametrize("M,Kdim", [(1, 256), (3, 1024), (8, 4096)])
def test_quant_q80_matches_numpy(M, Kdim):
    k = K()
    rng = np.random.default_rng(5)
    x = (rng.standard_normal((M, Kdim), dtype=np.float32) * 3).astype(np.float32)
    x[0, :32] = 0.0                                   # an all-zero block (d == 0)
    xq = k.alloc_xq(Kdim, 8, "cuda")
    xh = k.alloc_xh(Kdim, 16, "cuda")
    k.quant_q80(torch.from_numpy(x).cuda(), M, Kdim, 8, xq, None)
    k.quant_q80(torch.from_numpy(x).cuda(), M, Kdim, 16, None, xh)
    torch.cuda.synchronize()
    blocks = quant.quantize_q80(x)                    # numpy oracle, half-even
    ref = quant.dequantize_q80(blocks).reshape(M, Kdim)
    got, q, d, s = _decode_xq(xq, Kdim, 8, M)
    np.testing.assert_array_equal(got, ref)
    qref = blocks[:, 2:].copy().view(np.int8).reshape(M, Kdim // 32, 32).astype(np.int32)
    sums = s.transpose(1, 0, 2).reshape(M, Kdim // 32)
    np.testing.assert_array_equal(sums, 8 * qref.sum(-1))
    got_h = _decode_xh(xh, Kdim, 16, M)
    np.testing.assert_allclose(got_h, ref, rtol=1e-3, atol=1e-6)    # fp16 rounding of q*d


def test_xh_saturates():
    """fp16 activations of the tensor-core path: a block beyond the fp16 range saturates (no inf), everything else is
    the plain Q80 round trip."""
    k = K()
    x = np.zeros((2, 256), dtype=np.float32)
    x[0, :32] = np.linspace(-3e5, 3e5, 32)          # block scale 2362: far beyond fp16
    x[0, 32:64] = np.linspace(-100, 100, 32)
    x[1] = np.random.default_rng(3).standard_normal(256)
    xh = k.alloc_xh(256, 16, "cuda")
    k.quant_q80(torch.from_numpy(x).cuda(), 2, 256, 16, None, xh)
    torch.cuda.synchronize()
    got = _decode_xh(xh, 256, 16, 2)
    assert np.isfinite(got).all() and np.abs(got).max() <= 65504
    assert abs(got[0, 31]) > 6e4 and np.sign(got[0, 0]) == -1
    ref = quant.q80_roundtrip(x)
    np.testing.assert_allclose(got[0, 32:], ref[0, 32:], rtol=1e-3, atol=1e-6)
    np.testing.assert_allclose(got[1], ref[1], rtol=1e-3, atol=1e-6)


def _gemv_case(d, n, M, R, S, q80rt=False, seed=0):
    k = K()
    raw, wdq = _rand_q40(d, n, seed)
    w = k.make_q40t(torch.from_numpy(raw).cuda(), d, n)
    rng = np.random.default_rng(seed + 1)
    x = rng.standard_normal((M, n), dtype=np.float32)
    xq = k.alloc_xq(n, 8, "cuda")
    k.quant_q80(torch.from_numpy(x).cuda(), M, n, M, xq, None)
    xdq = quant.q80_roundtrip(x)
    ref = xdq.astype(np.float64) @ wdq.astype(np.float64).T
    y = torch.full((M, d), float("nan"), device="cuda")
    epi = k.epi_store([y.data_ptr()], d, 0, q80rt)
    k.gemv_q40(w, d, n, xq, M, epi, R, S)
    torch.cuda.synchronize()
    return y.cpu().numpy(), ref


@pytest.mark.parametrize("d,n,M,R,S", [
    (128, 128, 1, 1, 1), (256, 512, 1, 2, 1), (256, 1024, 2, 4, 2), (384, 2048, 4, 1, 4),
    (200, 1024, 8, 2, 8), (1024, 4096, 1, 1, 1), (512, 14336, 3, 4, 8), (128, 2304, 5, 1, 2),
])
def test_gemv_q40_matches_reference(d, n, M, R, S):
    got, ref = _gemv_case(d, n, M, R, S, seed=d + n)
    assert not np.isnan(got).any()
    scale = np.abs(ref).max()
    np.testing.assert_allclose(got, ref, rtol=2e-4, atol=2e-5 * scale)


def test_gemv_q80_roundtrip_epilogue():
    got, ref = _gemv_case(256, 512, 2, 1, 1, q80rt=True)
    ref_rt = quant.q80_roundtrip(ref.astype(np.float32))
    scale = np.abs(ref).max()
    np.testing.assert_allclose(got, ref_rt, rtol=0, atol=scale / 127.0 * 1.01)


def _gemm_case(d, n, M, S, seed=0, ncap_override=0):
    k = K()
    raw, wdq = _rand_q40(d, n, seed)
    w = k.make_q40t(torch.from_numpy(raw).cuda(), d, n)
    rng = np.random.default_rng(seed + 1)
    x = rng.standard_normal((M, n), dtype=np.float32)
    ncap = 16 if M <= 16 else 32 if M <= 32 else 64 if M <= 64 else (M + 127) // 128 * 128
    if ncap_override:
        ncap = ncap_override
    xh = k.alloc_xh(n, ncap, "cuda")
    k.quant_q80(torch.from_numpy(x).cuda(), M, n, ncap, None, xh)
    ref = quant.q80_roundtrip(x).astype(np.float64) @ wdq.astype(np.float64).T
    y = torch.full((M, d), float("nan"), device="cuda")
    epi = k.epi_store([y.data_ptr()], d, 0, False)
    k.gemm_q40_tc(w, d, n, xh, ncap, M, epi, S)
    torch.cuda.synchronize()
    return y.cpu().numpy(), ref


@pytest.mark.parametrize("d,n,M,S", [
    (128, 128, 16, 1), (128, 256, 9, 1), (256, 1024, 32, 1), (256, 1024, 32, 2), (384, 2048, 20, 4),
    (128, 4096, 64, 1), (512, 1024, 128, 1), (256, 512, 200, 1), (200, 1024, 40, 1),
    (256, 2048, 64, 2), (128, 2048, 128, 4), (256, 1024, 100, 8), (128, 1024, 256, 2), (128, 384, 300, 2),
])
def test_gemm_q40_tc_matches_reference(d, n, M, S):
    got, ref = _gemm_case(d, n, M, S, seed=d + n + M)
    assert not np.isnan(got).any()
    scale = np.abs(ref).max()
    # fp16 rounding of (nibble-8)*d and q*d: ~2^-11 relative per operand, random over K
    np.testing.assert_allclose(got, ref, rtol=2e-3, atol=2e-3 * scale)


def _decode_wh(wh: torch.Tensor, d: int, n: int) -> np.ndarray:
    """WH image -> [d_pad][n] fp16 values in natural K order (inverse of the layout in gemm_wh_tc.cu / xh_slot)."""
    rt, kt = (d + 127) // 128, n // 128
    a = wh.cpu().numpy().view(np.float16).reshape(rt, kt, 16, 16, 8, 8)        # [rt][kt][row group][chunk][row][slot]
    slot_of = np.zeros(32, dtype=np.int64)
    chunk_of = np.zeros(32, dtype=np.int64)
    for e in range(32):
        hi, e2 = e >> 4, e & 15
        w, r = e2 >> 2, e2 & 3
        chunk_of[e], slot_of[e] = w, hi * 4 + (2 if r == 1 else 1 if r == 2 else r)
    out = np.zeros((rt * 128, n), dtype=np.float32)
    for kb in range(4):
        for e in range(32):
            v = a[:, :, :, kb * 4 + chunk_of[e], :, slot_of[e]]                 # [rt][kt][16][8]
            out.reshape(rt, 128, kt, 128)[:, :, :, kb * 32 + e] = v.transpose(0, 2, 3, 1).reshape(rt, 128, kt)
    return out


def test_q40t_to_wh_matches_dequant():
    k = K()
    d, n = 300, 512
    raw, wdq = _rand_q40(d, n, 5)
    w = k.make_q40t(torch.from_numpy(raw).cuda(), d, n)
    wh = k.q40t_to_wh(w, d, n)
    torch.cuda.synchronize()
    got = _decode_wh(wh, d, n)
    np.testing.assert_array_equal(got[:d], wdq.astype(np.float16).astype(np.float32))
    assert not got[d:].any()


def _gemm_wh_case(d, n, M, ncap, force_cg, seed):
    k = K()
    raw, wdq = _rand_q40(d, n, seed)
    w = k.make_q40t(torch.from_numpy(raw).cuda(), d, n)
    wh = k.q40t_to_wh(w, d, n)
    rng = np.random.default_rng(seed + 1)
    x = rng.standard_normal((M, n), dtype=np.float32)
    xh = k.alloc_xh(n, ncap, "cuda")
    k.quant_q80(torch.from_numpy(x).cuda(), M, n, ncap, None, xh)
    ref = quant.q80_roundtrip(x).astype(np.float64) @ wdq.astype(np.float64).T
    y = torch.full((M, d), float("nan"), device="cuda")
    k.gemm_wh_tc(wh, d, n, xh, ncap, M, k.epi_store([y.data_ptr()], d, 0, False), force_cg)
    torch.cuda.synchronize()
    return y.cpu().numpy(), ref


@pytest.mark.parametrize("d,n,M,ncap,force_cg", [
    (256, 256, 256, 256, 1), (256, 256, 256, 256, 0), (128, 512, 40, 64, 0), (512, 1024, 300, 384, 0),
    (512, 1024, 300, 384, 1), (384, 384, 700, 768, 0), (200, 256, 256, 256, 0),
    # more tiles than CTAs / clusters: the persistent loop, both accumulators, barrier phases wrapping several times
    (9728, 1024, 1100, 1152, 0), (9600, 512, 520, 640, 0), (38912, 256, 260, 384, 0), (19456, 128, 512, 512, 1),
])
def test_gemm_wh_tc_matches_reference(d, n, M, ncap, force_cg):
    """Prompt-sized GEMM on pre-expanded weights: CTA pairs (cta_group::2) and single CTAs, partial activation tiles,
    padded weight rows, several tiles per CTA."""
    got, ref = _gemm_wh_case(d, n, M, ncap, force_cg, seed=d + n + M)
    assert not np.isnan(got).any()
    scale = np.abs(ref).max()
    np.testing.assert_allclose(got, ref, rtol=2e-3, atol=2e-3 * scale)


def _tc_tile(d, M, S, ncap, n_sms=148):
    """The activation tile width dl_gemm_q40_tc picks (mirror of gemm_q40_tc.cu:388-396), so the wide-tile tests
    below fail loudly if a change of the heuristic stops exercising the plans they are meant to cover."""
    row_tiles = (d + 127) // 128
    nt = 16 if M <= 16 else 32 if M <= 32 else 64 if M <= 64 else 128
    if M > 128 and ncap % 256 == 0 and row_tiles * ((M + 255) // 256) * S >= n_sms // 2:
        nt = 256
    while nt > 32 and row_tiles * ((M + nt - 1) // nt) * S < n_sms // 2:
        nt >>= 1
    return nt


@pytest.mark.parametrize("d,n,M,S,ncap,want_nt", [
    (9600, 1024, 64, 1, 64, 64), (9600, 1024, 50, 1, 64, 64), (9600, 1024, 128, 1, 128, 128), (9600, 1024, 100, 2, 128, 128),
    (9600, 1024, 256, 1, 256, 256), (9600, 1024, 300, 1, 512, 256), (4864, 1024, 256, 2, 256, 256), (9600, 512, 200, 1, 256, 256),
])
def test_gemm_q40_tc_wide_tiles(d, n, M, S, ncap, want_nt):
    """The prefill configurations: 64-, 128- and 256-wide activation tiles (different stage counts, accumulator
    counts, TMEM column maps and the chunked 256-wide epilogue) at shapes that really select them."""
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    assert _tc_tile(d, M, S, ncap, sms) == want_nt, "the tile heuristic changed: pick shapes that select this tile again"
    got, ref = _gemm_case(d, n, M, S, seed=d + n + M, ncap_override=ncap)
    assert not np.isnan(got).any()
    scale = np.abs(ref).max()
    np.testing.assert_allclose(got, ref, rtol=2e-3, atol=2e-3 * scale)


@pytest.mark.parametrize("M,ncap,want_nt", [(64, 64, 64), (128, 128, 128), (256, 256, 256)])
def test_wide_tile_epilogues(M, ncap, want_nt):
    """SiLU*mul -> Q80 slabs and RoPE + q / paged-KV write through the wide tensor-core tiles (the TTFT path)."""
    k = K()
    from dllama_b200.formats.mfile import LlmHeader
    from dllama_b200.models.oracle import apply_rope, rope_frequencies
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    dim = 512
    rng = np.random.default_rng(100 + M)
    x = rng.standard_normal((M, dim), dtype=np.float32) * 2
    xdq = quant.q80_roundtrip(x)
    xh = k.alloc_xh(dim, ncap, "cuda")
    k.quant_q80(torch.from_numpy(x).cuda(), M, dim, ncap, None, xh)
    # ---- SiLU * mul: 2 * hid rows interleaved (gate, up) ----
    hid = 4864                                       # 2 * hid = 76 row tiles; a multiple of 128 (XH slabs are per k-tile)
    assert _tc_tile(2 * hid, M, 1, ncap, sms) == want_nt
    raw1, w1 = _rand_q40(hid, dim, 31)
    raw3, w3 = _rand_q40(hid, dim, 32)
    w = torch.zeros(k.q40t_bytes(2 * hid, dim), dtype=torch.uint8, device="cuda")
    k.q40_repack(torch.from_numpy(raw1).cuda(), hid, dim, w, 2 * hid, 0, 2)
    k.q40_repack(torch.from_numpy(raw3).cuda(), hid, dim, w, 2 * hid, 1, 2)
    g, u = xdq @ w1.T, xdq @ w3.T
    ref = quant.q80_roundtrip((g / (1 + np.exp(-g)) * u).astype(np.float32))
    out = k.alloc_xh(hid, ncap, "cuda")
    k.gemm_q40_tc(w, 2 * hid, dim, xh, ncap, M, k.epi_silu_mul(None, out, ncap), 1)
    torch.cuda.synchronize()
    got = _decode_xh(out, hid, ncap, M)
    amax = np.abs(ref).max()
    np.testing.assert_allclose(got, ref, rtol=0, atol=amax / 127 * 1.01 + 5e-3 * amax)
    # ---- RoPE + q store + paged KV write ----
    hs, n_heads, n_kv = 128, 59, 8                  # q_dim 7552 + 2 * 1024 = 9600 rows: 75 row tiles
    q_dim, kv_dim = n_heads * hs, n_kv * hs
    d = q_dim + 2 * kv_dim
    assert _tc_tile(d, M, 1, ncap, sms) == want_nt
    h = LlmHeader(dim=q_dim, n_heads=n_heads, n_kv_heads=n_kv, rope_theta=500000.0, rope_type=2, rope_scaling_factor=8.0,
                  rope_scaling_low_freq_factor=1.0, rope_scaling_high_freq_factor=4.0, rope_scaling_orig_max_seq_len=8192)
    freqs = rope_frequencies(h)
    raw, wdq = _rand_q40(d, dim, 33)
    wq = k.make_q40t(torch.from_numpy(raw).cuda(), d, dim)
    y = torch.from_numpy(xdq @ wdq.T)
    positions = (np.arange(M) * 3 + 5).tolist()
    pos = torch.tensor(positions, dtype=torch.int32)
    q_ref = apply_rope(y[:, :q_dim], pos, freqs, hs)
    k_ref = apply_rope(y[:, q_dim:q_dim + kv_dim], pos, freqs, hs)
    v_ref = y[:, q_dim + kv_dim:]
    max_pages = (max(positions) + 32) // 32
    bt = torch.arange(max_pages, dtype=torch.int32).flip(0).reshape(1, max_pages).contiguous()        # one sequence, pages reversed
    Kc = torch.zeros(max_pages, n_kv, 32, hs, dtype=torch.float16, device="cuda")
    Vc = torch.zeros_like(Kc)
    q_out = torch.zeros(M, q_dim, device="cuda")
    epi = k.epi_qkv_rope(q_out, q_dim, kv_dim, hs, freqs.cuda(), pos.cuda(), torch.zeros(M, dtype=torch.int32, device="cuda"),
                         bt.cuda(), max_pages, n_kv, Kc, Vc)
    k.gemm_q40_tc(wq, d, dim, xh, ncap, M, epi, 1)
    torch.cuda.synchronize()
    sc = float(np.abs(q_ref.numpy()).max())
    np.testing.assert_allclose(q_out.cpu().numpy(), q_ref.numpy(), rtol=4e-3, atol=4e-3 * sc)
    for m in (0, 1, M // 2, M - 1):
        p = positions[m]
        page = int(bt[0, p // 32])
        np.testing.assert_allclose(Kc[page, :, p % 32, :].float().cpu().numpy().reshape(-1), k_ref[m].numpy(), rtol=6e-3, atol=6e-3 * sc)
        np.testing.assert_allclose(Vc[page, :, p % 32, :].float().cpu().numpy().reshape(-1), v_ref[m].numpy(), rtol=6e-3, atol=6e-3 * sc)


@pytest.mark.parametrize("d,n,M,acc", [(64, 256, 1, False), (100, 512, 3, False), (37, 1024, 8, True), (256, 4096, 20, False),
                                       (128, 128, 5, True)])
def test_gemm_f32_matches_reference(d, n, M, acc):
    """f32-weights path: plain f32 FMA matmul, incl. a column slice of a wider matrix (row stride > n)."""
    k = K()
    rng = np.random.default_rng(d + n + M)
    w_full = torch.from_numpy(rng.standard_normal((d, 2 * n), dtype=np.float32)).cuda()
    w = w_full[:, n:]                                   # column slice, like the wo / w2 slices of a TP rank
    x = torch.from_numpy(rng.standard_normal((M, n), dtype=np.float32)).cuda()
    y0 = torch.from_numpy(rng.standard_normal((M, d), dtype=np.float32)).cuda()
    y = y0.clone()
    out = k.gemm_f32(w, x, y, accumulate=acc)
    torch.cuda.synchronize()
    ref = x.double().cpu().numpy() @ w.double().cpu().numpy().T + (y0.double().cpu().numpy() if acc else 0.0)
    np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=2e-5, atol=2e-5 * np.abs(ref).max())


def test_device_sampler_distribution_and_determinism():
    """sample_kernel: argmax at temperature 0; with top-p every draw lies inside the nucleus the host sampler
    would build (sorted cumulative > topp) and frequencies follow the renormalised probabilities."""
    k = K()
    rng = np.random.default_rng(11)
    V, M = 3000, 4096
    base = (rng.standard_normal(V) * 2.5).astype(np.float32)
    logits = torch.from_numpy(np.tile(base, (M, 1))).cuda()
    coins = torch.from_numpy(((np.arange(M) + 0.5) / M).astype(np.float32)).cuda()       # stratified coins
    out = torch.zeros(M, dtype=torch.int32, device="cuda")
    temp, topp = 0.8, 0.9
    p = np.exp((base.astype(np.float64) - base.max()) / temp)
    p /= p.sum()
    order = np.argsort(-p)
    cut = int(np.searchsorted(np.cumsum(p[order]), topp)) + 1                             # crossing token included
    nucleus = set(order[:cut].tolist())
    k.sample_rows(logits, torch.full((M,), temp, device="cuda"), torch.full((M,), topp, device="cuda"), coins, out)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert set(got.tolist()) <= nucleus
    pn = np.zeros(V)
    pn[order[:cut]] = p[order[:cut]] / p[order[:cut]].sum()
    freq = np.bincount(got, minlength=V) / M
    assert np.abs(freq - pn).max() < 2.0 / M + 1e-3                                       # stratified coins: near-exact CDF inversion
    out2 = torch.zeros_like(out)
    k.sample_rows(logits, torch.full((M,), temp, device="cuda"), torch.full((M,), topp, device="cuda"), coins, out2)
    assert torch.equal(out, out2)                                                         # bit-deterministic
    # multinomial (top-p off) and greedy rows in one launch
    temps = torch.full((M,), 1.3, device="cuda")
    temps[::2] = 0.0
    k.sample_rows(logits, temps, torch.ones(M, device="cuda"), coins, out)
    got = out.cpu().numpy()
    assert (got[::2] == int(np.argmax(base))).all()
    p2 = np.exp((base.astype(np.float64) - base.max()) / 1.3)
    p2 /= p2.sum()
    freq2 = np.bincount(got[1::2], minlength=V) / (M // 2)
    assert np.abs(freq2 - p2).max() < 4.0 / M + 2e-3


def _attention_ref(q, kc, vc, pos, n_heads, n_kv, hs):
    M = q.shape[0]
    out = np.zeros((M, n_heads * hs), dtype=np.float64)
    G = n_heads // n_kv
    for m in range(M):
        T = pos[m] + 1
        for h_ in range(n_heads):
            kk = kc[m][:T, h_ // G].astype(np.float64)
            vv = vc[m][:T, h_ // G].astype(np.float64)
            s = kk @ q[m, h_ * hs:(h_ + 1) * hs].astype(np.float64) / math.sqrt(hs)
            s = np.exp(s - s.max())
            out[m, h_ * hs:(h_ + 1) * hs] = (s / s.sum()) @ vv
    return out


@pytest.mark.parametrize("n_heads,n_kv,hs,positions,n_splits,kv_dtype", [
    (8, 4, 32, [0, 5, 31, 32, 100], 1, torch.float16),
    (4, 1, 128, [200, 17], 4, torch.float16),
    (8, 8, 64, [63, 64, 65], 2, torch.float32),
    (16, 2, 128, [300], 8, torch.float16),
    (16, 1, 128, [129, 1000], 32, torch.float16),
])
def test_attention_matches_reference(n_heads, n_kv, hs, positions, n_splits, kv_dtype):
    k = K()
    M = len(positions)
    max_pages = (max(positions) + 32) // 32 + 1
    n_pages = M * max_pages
    rng = np.random.default_rng(7)
    q = rng.standard_normal((M, n_heads * hs), dtype=np.float32)
    kc = [rng.standard_normal((max_pages * 32, n_kv, hs), dtype=np.float32) for _ in range(M)]
    vc = [rng.standard_normal((max_pages * 32, n_kv, hs), dtype=np.float32) for _ in range(M)]
    perm = rng.permutation(n_pages).astype(np.int32)            # scattered physical pages
    bt = perm.reshape(M, max_pages)
    Kc = torch.zeros(n_pages, n_kv, 32, hs, dtype=kv_dtype)
    Vc = torch.zeros(n_pages, n_kv, 32, hs, dtype=kv_dtype)
    for m in range(M):
        for p in range(max_pages):
            Kc[bt[m, p]] = torch.from_numpy(kc[m][p * 32:(p + 1) * 32].transpose(1, 0, 2).copy()).to(kv_dtype)
            Vc[bt[m, p]] = torch.from_numpy(vc[m][p * 32:(p + 1) * 32].transpose(1, 0, 2).copy()).to(kv_dtype)
    # poison the tokens after pos (they must be masked, not multiplied by zero)
    kc_r = [Kc[torch.from_numpy(bt[m].astype(np.int64))].float().numpy().transpose(0, 2, 1, 3).reshape(-1, n_kv, hs)
            for m in range(M)]
    vc_r = [Vc[torch.from_numpy(bt[m].astype(np.int64))].float().numpy().transpose(0, 2, 1, 3).reshape(-1, n_kv, hs)
            for m in range(M)]
    for m in range(M):
        pg, off = divmod(positions[m] + 1, 32)
        if off:
            Kc[bt[m, pg], :, off:] = float("nan")
            Vc[bt[m, pg], :, off:] = float("inf")
    Kc, Vc = Kc.cuda(), Vc.cuda()
    y = torch.full((M, n_heads * hs), float("nan"), device="cuda")
    scratch = torch.zeros(k.attention_scratch_bytes(M, n_heads, hs, max(n_splits, 1)), dtype=torch.uint8,
                          device="cuda")
    pos_t = torch.tensor(positions, dtype=torch.int32, device="cuda")
    slot_t = torch.arange(M, dtype=torch.int32, device="cuda")
    bt_t = torch.from_numpy(bt).cuda()
    xq = k.alloc_xq(n_heads * hs, 8, "cuda") if (n_heads * hs) % 128 == 0 and M <= 8 else None
    for _ in range(2):      # twice: the split counters must self-reset
        k.attention(torch.from_numpy(q).cuda(), M, n_heads, n_kv, hs, Kc, Vc, pos_t, slot_t, bt_t, max_pages,
                    n_splits, scratch, xq, None, 8, y)
    torch.cuda.synchronize()
    ref = _attention_ref(q, kc_r, vc_r, positions, n_heads, n_kv, hs)
    got = y.cpu().numpy()
    assert not np.isnan(got).any()
    np.testing.assert_allclose(got, ref, rtol=2e-3, atol=2e-3)
    if xq is not None:
        dq, _, _, _ = _decode_xq(xq, n_heads * hs, 8, M)
        np.testing.assert_allclose(dq, quant.q80_roundtrip(got), rtol=0, atol=np.abs(got).max() / 127 * 1.01)


@pytest.mark.parametrize("n_heads,n_kv,hs,segs,tc", [
    (4, 1, 128, [(0, 0, 16)], False), (8, 2, 128, [(0, 37, 100)], False), (4, 4, 64, [(3, 5, 70)], False),
    (2, 1, 128, [(62, 200, 130)], False), (2, 2, 64, [(31, 31, 33)], False), (4, 2, 128, [(2, 10, 40), (45, 0, 90)], False),
    # the tcgen05 / TMEM kernel: one tile, partial tiles, several KV steps (lazy rescaling, both score buffers, stage
    # reuse), a chunk that starts deep inside a sequence, two segments
    (4, 1, 128, [(0, 0, 16)], True), (8, 2, 128, [(0, 37, 100)], True), (2, 1, 128, [(62, 200, 130)], True),
    (4, 2, 128, [(2, 10, 40), (45, 0, 90)], True), (4, 4, 128, [(0, 0, 700)], True), (2, 1, 128, [(5, 900, 300)], True),
    (8, 2, 128, [(0, 0, 257), (260, 130, 384)], True),
])
def test_attention_prefill_matches_reference(n_heads, n_kv, hs, segs, tc):
    """Tensor-core prompt-chunk attention.  segs = [(first row, first position, rows)]: every segment is a run of
    consecutive positions of its own sequence; rows outside the segments must not be touched (the per-row kernel
    serves them).  Compared with the f64 reference."""
    k = K()
    M = max(b + n for b, _, n in segs) + 2
    max_pages = max(p0 + n for _, p0, n in segs) // 32 + 2
    n_slots = len(segs)
    rng = np.random.default_rng(M)
    q = rng.standard_normal((M, n_heads * hs), dtype=np.float32)
    bt = rng.permutation(n_slots * max_pages).astype(np.int32).reshape(n_slots, max_pages)
    Kc = torch.zeros(n_slots * max_pages, n_kv, 32, hs, dtype=torch.float16)
    Vc = torch.zeros_like(Kc)
    kc_r, vc_r = [], []
    for s_i, (b, p0, n) in enumerate(segs):
        kc = rng.standard_normal((max_pages * 32, n_kv, hs), dtype=np.float32)
        vc = rng.standard_normal((max_pages * 32, n_kv, hs), dtype=np.float32)
        for p in range(max_pages):
            Kc[bt[s_i, p]] = torch.from_numpy(kc[p * 32:(p + 1) * 32].transpose(1, 0, 2).copy()).half()
            Vc[bt[s_i, p]] = torch.from_numpy(vc[p * 32:(p + 1) * 32].transpose(1, 0, 2).copy()).half()
        idx = torch.from_numpy(bt[s_i].astype(np.int64))
        kc_r.append(Kc[idx].float().numpy().transpose(0, 2, 1, 3).reshape(-1, n_kv, hs))
        vc_r.append(Vc[idx].float().numpy().transpose(0, 2, 1, 3).reshape(-1, n_kv, hs))
        pg, off = divmod(p0 + n, 32)                     # poison what lies after the last position
        if off:
            Kc[bt[s_i, pg], :, off:] = float("nan")
            Vc[bt[s_i, pg], :, off:] = float("inf")
    Kc, Vc = Kc.cuda(), Vc.cuda()
    positions, slots, meta = [0] * M, [0] * M, [len(segs)]
    for s_i, (b, p0, n) in enumerate(segs):
        for i in range(n):
            positions[b + i], slots[b + i] = p0 + i, s_i
        meta += [b, b + n]
    pos_t = torch.tensor(positions, dtype=torch.int32, device="cuda")
    slot_t = torch.tensor(slots, dtype=torch.int32, device="cuda")
    meta_t = torch.tensor(meta + [0] * 8, dtype=torch.int32, device="cuda")
    bt_t = torch.from_numpy(bt).cuda()
    y = torch.full((M, n_heads * hs), float("nan"), device="cuda")
    ncap = (M + 127) // 128 * 128
    xh = k.alloc_xh(n_heads * hs, ncap, "cuda") if (n_heads * hs) % 128 == 0 else None
    k.attention_prefill(torch.from_numpy(q).cuda(), M, n_heads, n_kv, hs, Kc, Vc, pos_t, slot_t, bt_t, max_pages, meta_t,
                        None, xh, ncap, y, tc=tc)
    torch.cuda.synchronize()
    got = y.cpu().numpy()
    covered = np.zeros(M, dtype=bool)
    for s_i, (b, p0, n) in enumerate(segs):
        covered[b:b + n] = True
        ref = _attention_ref(q[b:b + n], [kc_r[s_i]] * n, [vc_r[s_i]] * n, positions[b:b + n], n_heads, n_kv, hs)
        assert not np.isnan(got[b:b + n]).any()
        # fp16 Q / P operands: ~1e-3 relative
        np.testing.assert_allclose(got[b:b + n], ref, rtol=0, atol=4e-3 * max(1.0, np.abs(ref).max()))
    assert np.isnan(got[~covered]).all()                 # rows outside the segments are left alone


def test_qkv_rope_epilogue_and_kv_write():
    k = K()
    from dllama_b200.formats.mfile import LlmHeader
    from dllama_b200.models.oracle import apply_rope, rope_frequencies
    hs, n_heads, n_kv, dim, M = 64, 4, 2, 256, 3
    q_dim, kv_dim = n_heads * hs, n_kv * hs
    d = q_dim + 2 * kv_dim
    h = LlmHeader(dim=q_dim, n_heads=n_heads, n_kv_heads=n_kv, rope_theta=500000.0, rope_type=2,
                  rope_scaling_factor=32.0, rope_scaling_low_freq_factor=1.0, rope_scaling_high_freq_factor=4.0,
                  rope_scaling_orig_max_seq_len=8192)
    freqs = rope_frequencies(h)
    raw, wdq = _rand_q40(d, dim, 11)
    w = k.make_q40t(torch.from_numpy(raw).cuda(), d, dim)
    rng = np.random.default_rng(12)
    x = rng.standard_normal((M, dim), dtype=np.float32)
    xq = k.alloc_xq(dim, 8, "cuda")
    k.quant_q80(torch.from_numpy(x).cuda(), M, dim, M, xq, None)
    y = torch.from_numpy(quant.q80_roundtrip(x) @ wdq.T)
    positions = [0, 37, 70]
    pos = torch.tensor(positions, dtype=torch.int32)
    q_ref = apply_rope(y[:, :q_dim], pos, freqs, hs)
    k_ref = apply_rope(y[:, q_dim:q_dim + kv_dim], pos, freqs, hs)
    v_ref = y[:, q_dim + kv_dim:]
    max_pages, n_pages = 4, 12
    bt = torch.tensor([[3, 7, 1, 0], [2, 5, 9, 4], [11, 10, 8, 6]], dtype=torch.int32)
    for kv_dtype in (torch.float16, torch.float32):
        Kc = torch.zeros(n_pages, n_kv, 32, hs, dtype=kv_dtype, device="cuda")
        Vc = torch.zeros_like(Kc)
        q_out = torch.zeros(M, q_dim, device="cuda")
        epi = k.epi_qkv_rope(q_out, q_dim, kv_dim, hs, freqs.cuda(), pos.cuda(),
                             torch.arange(M, dtype=torch.int32, device="cuda"), bt.cuda(), max_pages, n_kv, Kc, Vc)
        k.gemv_q40(w, d, dim, xq, M, epi, 2, 2)
        torch.cuda.synchronize()
        np.testing.assert_allclose(q_out.cpu().numpy(), q_ref.numpy(), rtol=1e-3, atol=2e-4)
        tol = 2e-3 if kv_dtype == torch.float16 else 2e-4
        for m, p in enumerate(positions):
            page = int(bt[m, p // 32])
            got_k = Kc[page, :, p % 32, :].float().cpu().numpy().reshape(-1)
            got_v = Vc[page, :, p % 32, :].float().cpu().numpy().reshape(-1)
            np.testing.assert_allclose(got_k, k_ref[m].numpy(), rtol=tol, atol=tol)
            np.testing.assert_allclose(got_v, v_ref[m].numpy(), rtol=tol, atol=tol)
        assert float(Kc.abs().sum()) > 0 and int((Kc != 0).any(dim=-1).sum()) == M * n_kv


@pytest.mark.parametrize("M,use_tc", [(2, False), (24, True)])
def test_silu_mul_epilogue(M, use_tc):
    k = K()
    hid, dim = 256, 256
    raw1, w1 = _rand_q40(hid, dim, 21)
    raw3, w3 = _rand_q40(hid, dim, 22)
    w = torch.zeros(k.q40t_bytes(2 * hid, dim), dtype=torch.uint8, device="cuda")
    k.q40_repack(torch.from_numpy(raw1).cuda(), hid, dim, w, 2 * hid, 0, 2)
    k.q40_repack(torch.from_numpy(raw3).cuda(), hid, dim, w, 2 * hid, 1, 2)
    rng = np.random.default_rng(23)
    x = rng.standard_normal((M, dim), dtype=np.float32) * 4
    xdq = quant.q80_roundtrip(x)
    g, u = xdq @ w1.T, xdq @ w3.T
    ref = quant.q80_roundtrip((g / (1 + np.exp(-g)) * u).astype(np.float32))
    if use_tc:
        xh = k.alloc_xh(dim, 32, "cuda")
        k.quant_q80(torch.from_numpy(x).cuda(), M, dim, 32, None, xh)
        out = k.alloc_xh(hid, 32, "cuda")
        k.gemm_q40_tc(w, 2 * hid, dim, xh, 32, M, k.epi_silu_mul(None, out, 32), 1)
        torch.cuda.synchronize()
        got = _decode_xh(out, hid, 32, M)
        tol = 5e-3
    else:
        xq = k.alloc_xq(dim, 8, "cuda")
        k.quant_q80(torch.from_numpy(x).cuda(), M, dim, M, xq, None)
        out = k.alloc_xq(hid, M, "cuda")
        k.gemv_q40(w, 2 * hid, dim, xq, M, k.epi_silu_mul(out, None, M), 2, 2)
        torch.cuda.synchronize()
        got, _, _, _ = _decode_xq(out, hid, M, M)
        tol = 1e-4
    amax = np.abs(ref).max()
    np.testing.assert_allclose(got, ref, rtol=0, atol=amax / 127 * 1.01 + tol * amax)


def test_reduce_norm_quant_with_partials():
    k = K()
    M, dim, n_src = 3, 512, 4
    rng = np.random.default_rng(31)
    x = rng.standard_normal((M, dim), dtype=np.float32)
    parts = rng.standard_normal((n_src, 8, dim), dtype=np.float32)
    wn = rng.standard_normal(dim).astype(np.float32)
    xin = torch.from_numpy(x).cuda()
    xout = torch.zeros_like(xin)
    pt = torch.from_numpy(parts).cuda()
    yout = torch.zeros(M, dim, device="cuda")
    xq = k.alloc_xq(dim, 8, "cuda")
    k.reduce_norm_quant(xin, xout, pt.data_ptr(), n_src, 8 * dim, None, None, 0, 0, torch.from_numpy(wn).cuda(),
                        1e-5, dim, M, 8, xq, None, yout)
    torch.cuda.synchronize()
    xs = x + parts[:, :M].sum(0)
    np.testing.assert_allclose(xout.cpu().numpy(), xs, rtol=1e-6, atol=1e-6)
    yref = wn * (xs / np.sqrt((xs * xs).mean(-1, keepdims=True) + 1e-5))
    np.testing.assert_allclose(yout.cpu().numpy(), yref, rtol=1e-5, atol=1e-5)
    got, _, _, _ = _decode_xq(xq, dim, 8, M)
    np.testing.assert_allclose(got, quant.q80_roundtrip(yout.cpu().numpy()), rtol=0, atol=1e-6)


def test_q80_wire_rows_producer_to_consumer():
    """Tensor-parallel hand-off of tensor-core batches under the flag protocol: the matmul epilogue pushes its partial
    sums as Q80 wire rows (ld int8 + ld/32 fp16 scales per row) into N destination buffers (standing in for the peers)
    and releases their flags; reduce_norm_quant waits for the flags and merges the rows.  Must equal the f32 exchange
    of the Q80 round trip bit for bit (q * d is exact in f32)."""
    k = K()
    M, d, n, n_src, rows_cap = 40, 512, 256, 2, 64
    rng = np.random.default_rng(71)
    step = torch.tensor([3], dtype=torch.int32, device="cuda")
    flag_mul, flag_add = 10, 4
    rowb = d + d // 16
    src_stride = rows_cap * d                                     # elements between sources in the f32 layout
    bufs = [torch.zeros(n_src * rows_cap * rowb + 64, dtype=torch.uint8, device="cuda") for _ in range(2)]
    flags = [torch.zeros(8, dtype=torch.int32, device="cuda") for _ in range(2)]
    counter = torch.zeros(1, dtype=torch.int32, device="cuda")
    ncap = 64
    parts_ref = []
    for src in range(n_src):                                      # source `src` writes its slot in both destinations
        raw, wdq = _rand_q40(d, n, 80 + src)
        w = k.make_q40t(torch.from_numpy(raw).cuda(), d, n)
        xa = rng.standard_normal((M, n), dtype=np.float32)
        xh = k.alloc_xh(n, ncap, "cuda")
        k.quant_q80(torch.from_numpy(xa).cuda(), M, n, ncap, None, xh)
        epi = k.epi_store([b.data_ptr() for b in bufs], d, src * src_stride, True, [f.data_ptr() + 4 * src for f in flags],
                          step.data_ptr(), flag_mul, flag_add, counter.data_ptr(), wire_q80=True)
        k.gemm_q40_tc(w, d, n, xh, ncap, M, epi, 1)
        y = torch.zeros(M, d, device="cuda")                      # the same product through the plain f32 epilogue
        k.gemm_q40_tc(w, d, n, xh, ncap, M, k.epi_store([y.data_ptr()], d, 0, True), 1)
        parts_ref.append(y)
    torch.cuda.synchronize()
    assert all(int(f[s]) == 3 * flag_mul + flag_add for f in flags for s in range(n_src))
    x = rng.standard_normal((M, d), dtype=np.float32)
    wn = (1.0 + 0.1 * rng.standard_normal(d)).astype(np.float32)
    outs = []
    for b, f in zip(bufs, flags):
        xout = torch.zeros(M, d, device="cuda")
        yout = torch.zeros(M, d, device="cuda")
        k.reduce_norm_quant(torch.from_numpy(x).cuda(), xout, b.data_ptr(), n_src, src_stride, f.data_ptr(), step.data_ptr(),
                            flag_mul, flag_add, torch.from_numpy(wn).cuda(), 1e-5, d, M, ncap, None, k.alloc_xh(d, ncap, "cuda"),
                            yout, wire_q80=True)
        outs.append((xout, yout))
    torch.cuda.synchronize()
    xs = torch.from_numpy(x).cuda()
    for p in parts_ref:                                            # rank order, like the kernel
        xs = xs + p
    for xout, yout in outs:
        assert torch.equal(xout, xs)
        xr = xs.cpu().numpy()
        yref = wn * (xr / np.sqrt((xr * xr).mean(-1, keepdims=True) + 1e-5))
        np.testing.assert_allclose(yout.cpu().numpy(), yref, rtol=1e-5, atol=1e-5)


def test_ll_exchange_words_producer_to_consumers():
    """Tensor-parallel hand-off in LL words: a matmul epilogue pushes {value, epoch} pairs into N destination
    buffers (standing in for the peers), then the fused GEMV prologue and the stand-alone rnq kernel consume
    them by polling the epoch in the data -- no flags.  A stale epoch in the buffer must not be accepted."""
    k = K()
    M, d, n, n_src = 2, 512, 256, 2
    raw, wdq = _rand_q40(d, n, 61)
    w = k.make_q40t(torch.from_numpy(raw).cuda(), d, n)
    rng = np.random.default_rng(62)
    xa = rng.standard_normal((M, n), dtype=np.float32)
    xq = k.alloc_xq(n, M, "cuda")
    k.quant_q80(torch.from_numpy(xa).cuda(), M, n, M, xq, None)
    step = torch.tensor([7], dtype=torch.int32, device="cuda")
    flag_mul, flag_add = 33, 5
    epoch = 7 * flag_mul + flag_add
    src_stride = 8 * d                                   # elements between sources (row capacity 8)
    # every "rank" holds [n_src][8][d] LL words; pre-fill with a stale epoch and a poison value
    bufs = [torch.zeros(n_src * src_stride, 2, dtype=torch.int32, device="cuda") for _ in range(2)]
    for b in bufs:
        b[:, 0] = 0x7FC00000                              # NaN bits
        b[:, 1] = epoch - 1
    ref_part = quant.q80_roundtrip((quant.q80_roundtrip(xa).astype(np.float64) @ wdq.astype(np.float64).T).astype(np.float32))
    for src in range(n_src):                              # source `src` writes its slot in both destinations
        epi = k.epi_store([b.data_ptr() for b in bufs], d, src * src_stride, True, None, step.data_ptr(), flag_mul, flag_add,
                          ll=True)
        k.gemv_q40(w, d, n, xq, M, epi, 1, 2)
    torch.cuda.synchronize()
    words = bufs[1].cpu().numpy().reshape(n_src, 8, d, 2)
    assert (words[:, :M, :, 1] == epoch).all() and (words[:, M:, :, 1] == epoch - 1).all()
    got_part = words[:, :M, :, 0].copy().view(np.float32)
    np.testing.assert_allclose(got_part[0], ref_part, rtol=0, atol=np.abs(ref_part).max() / 127 * 1.01)
    np.testing.assert_array_equal(got_part[0], got_part[1])
    # consumers: x_out = x_in + sum of the sources
    x = rng.standard_normal((8, d), dtype=np.float32)
    wn = (1.0 + 0.1 * rng.standard_normal(d)).astype(np.float32)
    xin = torch.from_numpy(x).cuda()
    xs = x[:M] + got_part.sum(0)
    xout = torch.full_like(xin, float("nan"))
    yout = torch.zeros(M, d, device="cuda")
    k.reduce_norm_quant(xin, xout, bufs[0].data_ptr(), n_src, src_stride, None, step.data_ptr(), flag_mul, flag_add,
                        torch.from_numpy(wn).cuda(), 1e-5, d, M, 8, None, None, yout, ll=True)
    torch.cuda.synchronize()
    np.testing.assert_allclose(xout.cpu().numpy()[:M], xs, rtol=1e-6, atol=1e-6)
    raw2, wdq2 = _rand_q40(128, d, 63)
    w2 = k.make_q40t(torch.from_numpy(raw2).cuda(), 128, d)
    xout2 = torch.full_like(xin, float("nan"))
    y = torch.full((M, 128), float("nan"), device="cuda")
    nm = k.norm_prolog(xin, xout2, bufs[1].data_ptr(), n_src, src_stride, None, step.data_ptr(), flag_mul, flag_add,
                       torch.from_numpy(wn).cuda(), 1e-5, ll=True)
    k.gemv_q40(w2, 128, d, None, M, k.epi_store([y.data_ptr()], 128), 1, 2, norm=nm)
    torch.cuda.synchronize()
    np.testing.assert_allclose(xout2.cpu().numpy()[:M], xs, rtol=1e-6, atol=1e-6)
    yn = wn * (xs / np.sqrt((xs * xs).mean(-1, keepdims=True) + 1e-5))
    ref = quant.q80_roundtrip(yn.astype(np.float32)).astype(np.float64) @ wdq2.astype(np.float64).T
    np.testing.assert_allclose(y.cpu().numpy(), ref, rtol=2e-3, atol=2e-3 * np.abs(ref).max())


def test_argmax_and_embedding():
    k = K()
    logits = torch.randn(5, 1000, device="cuda")
    logits[2, 17] = 100.0
    logits[3, :] = 0.0                 # ties -> lowest index
    out = torch.zeros(5, dtype=torch.int32, device="cuda")
    k.argmax(logits, 5, 1000, 1000, out)
    torch.cuda.synchronize()
    ref = logits.argmax(dim=-1).cpu().numpy()
    got = out.cpu().numpy()
    assert got[2] == 17 and got[3] == 0
    np.testing.assert_array_equal(got[[0, 1, 4]], ref[[0, 1, 4]])
    big = torch.randn(600, 777, device="cuda")          # more rows than the scratch holds at once: 256-row passes
    out = torch.zeros(600, dtype=torch.int32, device="cuda")
    k.argmax(big, 600, 777, 777, out)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(out.cpu().numpy(), big.argmax(dim=-1).cpu().numpy())
    table = torch.randn(50, 256, device="cuda")
    toks = torch.tensor([3, 49, 0], dtype=torch.int32, device="cuda")
    x = torch.zeros(3, 256, device="cuda")
    k.embedding(table, toks, x, 3, 256, None)
    torch.cuda.synchronize()
    torch.testing.assert_close(x, table[[3, 49, 0]])


@pytest.mark.parametrize("M,d,dim,n_src,R,S", [(1, 256, 512, 0, 1, 1), (1, 384, 1024, 4, 1, 4), (3, 256, 2048, 8, 2, 8),
                                                (8, 128, 4096, 2, 1, 2), (2, 512, 4096, 1, 4, 1)])
def test_gemv_fused_norm_prologue(M, d, dim, n_src, R, S):
    """merge-add + RMSNorm + Q80 computed inside the GEMV (cluster ranks share the sum of squares)."""
    k = K()
    raw, wdq = _rand_q40(d, dim, 41)
    w = k.make_q40t(torch.from_numpy(raw).cuda(), d, dim)
    rng = np.random.default_rng(42)
    x = rng.standard_normal((8, dim), dtype=np.float32)
    parts = rng.standard_normal((max(n_src, 1), 8, dim), dtype=np.float32) * 0.5
    wn = (1.0 + 0.1 * rng.standard_normal(dim)).astype(np.float32)
    xin = torch.from_numpy(x).cuda()
    xout = torch.full_like(xin, float("nan"))
    pt = torch.from_numpy(parts).cuda()
    y = torch.full((M, d), float("nan"), device="cuda")
    nm = k.norm_prolog(xin, xout, pt.data_ptr() if n_src else None, n_src, 8 * dim, None, None, 0, 0,
                       torch.from_numpy(wn).cuda(), 1e-5)
    k.gemv_q40(w, d, dim, None, M, k.epi_store([y.data_ptr()], d), R, S, norm=nm)
    torch.cuda.synchronize()
    xs = x[:M].copy()
    for s in range(n_src):
        xs = xs + parts[s, :M]
    np.testing.assert_allclose(xout.cpu().numpy()[:M], xs, rtol=1e-6, atol=1e-6)
    yn = wn * (xs / np.sqrt((xs * xs).mean(-1, keepdims=True) + 1e-5))
    ref = quant.q80_roundtrip(yn.astype(np.float32)).astype(np.float64) @ wdq.astype(np.float64).T
    got = y.cpu().numpy()
    assert not np.isnan(got).any()
    # a Q80 rounding tie can flip when the f32 sum order differs: allow one quantum on one element
    np.testing.assert_allclose(got, ref, rtol=2e-3, atol=2e-3 * np.abs(ref).max())


def test_gemv_residual_epilogue_in_place():
    k = K()
    d, n, M = 256, 512, 2
    raw, wdq = _rand_q40(d, n, 51)
    w = k.make_q40t(torch.from_numpy(raw).cuda(), d, n)
    rng = np.random.default_rng(52)
    x = rng.standard_normal((M, n), dtype=np.float32)
    res = rng.standard_normal((M, d), dtype=np.float32)
    xq = k.alloc_xq(n, M, "cuda")
    k.quant_q80(torch.from_numpy(x).cuda(), M, n, M, xq, None)
    r = torch.from_numpy(res).cuda()
    k.gemv_q40(w, d, n, xq, M, k.epi_store([r.data_ptr()], d, 0, True, residual=r.data_ptr(), residual_ld=d), 1, 2)
    torch.cuda.synchronize()
    prod = quant.q80_roundtrip(x) @ wdq.T
    ref = res + quant.q80_roundtrip(prod.astype(np.float32))
    np.testing.assert_allclose(r.cpu().numpy(), ref, rtol=0, atol=np.abs(prod).max() / 127 * 1.01)


@pytest.mark.parametrize("d,n,M,R,S,T", [
    (1024, 512, 2, 1, 1, 3), (2048, 256, 1, 1, 1, 4), (640, 1024, 1, 1, 3, 1), (384, 2048, 3, 1, 5, 1),
    (896, 1536, 8, 1, 7, 1), (2176, 384, 4, 1, 1, 17), (256, 4096, 1, 2, 6, 1),
])
def test_gemv_tiles_per_cta_and_odd_clusters(d, n, M, R, S, T):
    """T row tiles streamed by one CTA (pipeline crosses tile boundaries); cluster sizes that are not
    powers of two (uneven K ranges)."""
    k = K()
    raw, wdq = _rand_q40(d, n, d + n + T)
    w = k.make_q40t(torch.from_numpy(raw).cuda(), d, n)
    rng = np.random.default_rng(T + S)
    x = rng.standard_normal((M, n), dtype=np.float32)
    xq = k.alloc_xq(n, M, "cuda")
    k.quant_q80(torch.from_numpy(x).cuda(), M, n, M, xq, None)
    ref = quant.q80_roundtrip(x).astype(np.float64) @ wdq.astype(np.float64).T
    y = torch.full((M, d), float("nan"), device="cuda")
    for _ in range(2):
        k.gemv_q40(w, d, n, xq, M, k.epi_store([y.data_ptr()], d), R, S, tiles_per_cta=T)
    torch.cuda.synchronize()
    got = y.cpu().numpy()
    assert not np.isnan(got).any()
    np.testing.assert_allclose(got, ref, rtol=2e-4, atol=2e-5 * np.abs(ref).max())


def test_choose_splits_stays_within_one_wave():
    k = K()
    from dllama_b200.ops.lib import EPI_QKV_ROPE, EPI_SILU_MUL, EPI_STORE
    for d, n, kind in [(6144, 4096, EPI_QKV_ROPE), (4096, 4096, EPI_STORE), (28672, 4096, EPI_SILU_MUL),
                       (4096, 14336, EPI_STORE), (128256, 4096, EPI_STORE), (768, 4096, EPI_QKV_ROPE),
                       (4096, 512, EPI_STORE), (3584, 4096, EPI_SILU_MUL), (4096, 1792, EPI_STORE), (16032, 4096, EPI_STORE)]:
        for M in (1, 8):
            R, S, T = k.choose_splits(d, n, kind, 148, M)
            units = ((d + 127) // 128) * R
            grid = ((units + T - 1) // T) * S
            assert grid <= 148 and (T == 1 or (R == 1 and S == 1)) and S <= n // 128
            assert k.cuda_lib().dl_gemv_q40_smem_bytes(n, M, S) <= 227 * 1024
