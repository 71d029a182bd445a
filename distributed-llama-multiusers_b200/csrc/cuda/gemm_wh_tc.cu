// Prompt-sized Q40 x Q80 matmul (M >= a few hundred activation rows) on tcgen05, two phases:
//
//   1. q40t_to_wh_kernel expands a Q40T weight image ONCE into "WH": fp16 tiles of 128 rows x 128 K in the UMMA
//      K-major no-swizzle core-matrix image (the same 8 rows x 16 B core matrices and the same K order inside a
//      32-block as the XH activation slabs).  fp16((nibble - 8) * d), one rounding: bit-identical to what the
//      on-the-fly dequantisers of gemm_q40_tc.cu feed the tensor core.  The engine keeps the WH copy resident when
//      it fits (8B: 15 GB of the 180 GB) or rebuilds it per matrix into a scratch buffer.
//   2. gemm_wh_tc_kernel is then a plain fp16 SS-mode GEMM: both operands arrive by bulk async copies (UBLKCP) of
//      contiguous, pre-tiled 16 KB pieces; nothing is unpacked inside the loop, so the unpack -- the limiter of the
//      fused kernel (ALU pipe 50 %, tensor pipe 14 % in profiles/ncu_tc_tmemA_w13_m32.txt), which every M tile
//      repeated -- is paid once per weight instead of once per 256 activation rows.
//
// gemm_wh_tc_kernel<2> pairs two CTAs (cluster of 2 = one TPC) on a 256 weight rows x 256 activation rows tile with
// `tcgen05.mma.cta_group::2`: each CTA stages its own 128 weight rows and HALF of the activation rows (64 KB per
// 128-wide k step and CTA instead of 96 KB), which is what keeps the L2 -> shared-memory feed (about 43 B/clk per SM)
// below the tensor pipe's appetite.  gemm_wh_tc_kernel<1> is the single-CTA variant (odd row-tile counts).
//
// Persistent CTAs walk the tile list (activation tiles fastest, so the CTAs that run at the same time share one
// weight panel in L2); each CTA owns 128 TMEM lanes x 2 accumulators of 256 columns, so the epilogue of tile i
// (TMEM -> registers -> smem chunk -> the shared fused epilogues of epilogue.cuh) overlaps the MMAs of tile i+1.
//
// Warp roles (576 threads): warp 0 producer (4-6 lanes issue one 16 KB bulk copy each per stage), warp 1 MMA issuer + TMEM owner (in the
// second CTA of a pair: forwards "my half of the stage has landed" to the leader's barrier), warps 2-17 epilogue.
// Reference semantics: matmul_Q80_Q40_F32 (src/nn/nn-cpu-ops.cpp:352-377), llamafile sgemm (src/nn/llamafile/sgemm.cpp:593-783).
#include "epilogue.cuh"

#define WH_TILE_BYTES 32768
#define WT_NT 256                      // activation rows per tile
#define WT_THREADS 576
#define WT_EPI_CHUNK 32

struct WhArgs {
    const uint8_t* wh;                 // [row_tiles][KT][32 KB]
    const uint8_t* xh;                 // [KT][ncap/8][2048]
    int d_real, row_tiles, KT, ncap, M, m_tiles, n_tiles;
    int n_epi_ctas;
    int piece;                         // bytes per bulk copy (4096 / 8192 / 16384)
    int pf_dist, pf_mask;              // weight tiles are pulled into L2 pf_dist k-steps ahead by the CTAs whose activation tile index & pf_mask == 0
    int dbg;                           // timing experiments (results invalid): 1 = no epilogue work, 2 = no MMAs, 4 = no activation copies
    unsigned long long* tr;
    DlEpilogue epi;
};

template <int CG> DL_DEVINL void wt_tmem_alloc(uint32_t* holder) {
    if (CG == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(holder)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(holder)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
}
template <int CG> DL_DEVINL void wt_tmem_dealloc(uint32_t taddr) {
    if (CG == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(512) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(512) : "memory");
}
template <int CG> DL_DEVINL void wt_mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    if (CG == 1) {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
    } else {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
    }
}
// all MMAs issued so far by this thread have completed -> arrive on the barrier (in both CTAs of a pair)
template <int CG> DL_DEVINL void wt_commit(uint64_t* bar) {
    if (CG == 1) {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
    } else {
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
                     "h"((uint16_t)3)
                     : "memory");
    }
}
DL_DEVINL void wt_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// Bounded waits: a protocol bug must end in a trap (an error the host sees), never in a hung GPU.
#define WT_DEADLINE_NS 4000000000ull
DL_DEVINL void wt_wait(uint64_t* bar, uint32_t parity, unsigned long long t0) {
    unsigned n = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++n & 0x3FFu) == 0 && globaltimer_ns() - t0 > WT_DEADLINE_NS) __trap();
    }
}
DL_DEVINL void wt_wait_cluster(uint64_t* bar, uint32_t parity, unsigned long long t0) {
    unsigned n = 0;
    uint32_t ok;
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (ok) break;
        if ((++n & 0x3FFu) == 0 && globaltimer_ns() - t0 > WT_DEADLINE_NS) __trap();
    }
}

template <int CG>
__global__ void __launch_bounds__(WT_THREADS, 1) gemm_wh_tc_kernel(const __grid_constant__ WhArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    constexpr int NS = (CG == 1) ? 2 : 3;                 // stages of one 128-wide k step
    constexpr int B_ROWS = WT_NT / CG;                    // activation rows staged by this CTA
    constexpr int STAGE = WH_TILE_BYTES + B_ROWS * 256;   // 96 KB / 64 KB
    trace_begin(a.tr);
    const unsigned long long t0 = globaltimer_ns();
    float* out_s = reinterpret_cast<float*>(smem + NS * STAGE);          // [32][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(out_s + WT_EPI_CHUNK * 128);
    uint64_t* full = bars;               // [NS]  this CTA's copies of a stage have landed
    uint64_t* peer_full = bars + 4;      // [NS]  (leader of a pair) the other CTA's copies have landed
    uint64_t* empty = bars + 8;          // [NS]  the MMAs reading the stage have retired (both CTAs)
    uint64_t* acc_full = bars + 12;      // [2]   accumulator complete
    uint64_t* acc_empty = bars + 14;     // [2]   accumulator drained by the epilogue(s)
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 16);
    int* sflag = reinterpret_cast<int*>(tmem_holder + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint32_t crank = 0;
    if (CG == 2) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
    const int cid = blockIdx.x / CG, ncl = gridDim.x / CG;
    const int KT = a.KT, m_tiles = a.m_tiles;

    if (tid == 0) {
        for (int s = 0; s < NS; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&peer_full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int u = 0; u < 2; u++) {
            mbar_init(&acc_full[u], 1);
            mbar_init(&acc_empty[u], CG);
        }
        mbar_fence_init();
    }
    if (warp == 1) wt_tmem_alloc<CG>(tmem_holder);
    tc_fence_before();
    __syncthreads();
    if (CG == 2) wt_cluster_sync();      // the partner's barriers exist before anything arrives on them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
    pdl_launch_dependents();

    if (warp == 0) {
        // ------------------------------------ producer ------------------------------------
        // the stage is fetched as `piece`-byte bulk copies, one per lane: first the weight tile, then row groups of the
        // activation slab (many requests in flight: one bulk copy streams at only ~16 B/clk)
        const int piece = a.piece, npiece = STAGE / piece, na = WH_TILE_BYTES / piece, prow = piece / 256;
        if (lane < npiece) {
            pdl_wait();
            const size_t x_kt_stride = (size_t)(a.ncap >> 3) * XH_ROWGROUP_BYTES;
            uint32_t it = 0;
            for (int t = cid; t < a.n_tiles; t += ncl) {
                const int rp = t / m_tiles, mt = t - rp * m_tiles;
                const int rt = rp * CG + (int)crank;
                const int mrow0 = mt * WT_NT + (int)crank * B_ROWS;
                const int b_rows = max(0, min(B_ROWS, a.ncap - mrow0));
                const uint8_t* src;
                size_t stride;
                uint32_t bytes;
                if (lane < na) {
                    src = a.wh + (size_t)rt * KT * WH_TILE_BYTES + lane * piece;
                    stride = WH_TILE_BYTES;
                    bytes = piece;
                } else {
                    const int r0 = (lane - na) * prow;
                    src = a.xh + (size_t)((mrow0 + r0) >> 3) * XH_ROWGROUP_BYTES;
                    stride = x_kt_stride;
                    bytes = (a.dbg & 4) ? 0u : (uint32_t)max(0, min(prow, b_rows - r0)) * 256u;
                }
                const uint32_t total = WH_TILE_BYTES + ((a.dbg & 4) ? 0u : (uint32_t)b_rows * 256u);
                const bool pf = lane < na && a.pf_dist > 0 && (mt & a.pf_mask) == 0;
                for (int kt = 0; kt < KT; kt++, it++) {
                    if (pf) {      // optional: pull the weight tile pf_dist k-steps ahead into L2 (measured: no gain, off by default)
                        int j = kt + a.pf_dist, tt = t;
                        while (j >= KT) { j -= KT; tt += ncl; }
                        if (tt < a.n_tiles) {
                            const int rt2 = (tt / m_tiles) * CG + (int)crank;
                            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.wh + ((size_t)rt2 * KT + j) * WH_TILE_BYTES + lane * piece),
                                         "r"(piece)
                                         : "memory");
                        }
                    }
                    const uint32_t s = it % NS;
                    if (it >= NS) wt_wait(&empty[s], ((it / NS) & 1) ^ 1, t0);
                    if (lane == 0) mbar_arrive_expect_tx(&full[s], total);
                    if (bytes) bulk_g2s(smem + s * STAGE + lane * piece, src + (size_t)kt * stride, bytes, &full[s]);
                }
            }
        }
    } else if (warp == 1) {
        if (CG == 1 || crank == 0) {
            // ------------------------------------ MMA issuer ------------------------------------
            const uint32_t idesc = umma_idesc_f16_f32(128 * CG, WT_NT);
            const uint32_t desc_hi = (2048u >> 4) | (1u << 14);      // SBO 2048 B (8-row groups), descriptor version 1
            uint32_t it = 0, tcount = 0;
            for (int t = cid; t < a.n_tiles; t += ncl, tcount++) {
                const uint32_t u = tcount & 1;
                if (tcount >= 2) {      // the epilogue(s) of the tile that last used this accumulator are done with it
                    if (CG == 2) wt_wait_cluster(&acc_empty[u], ((tcount >> 1) & 1) ^ 1, t0);
                    else wt_wait(&acc_empty[u], ((tcount >> 1) & 1) ^ 1, t0);
                }
                tc_fence_after();
                for (int kt = 0; kt < KT; kt++, it++) {
                    const uint32_t s = it % NS, ph = (it / NS) & 1;
                    wt_wait(&full[s], ph, t0);
                    if (CG == 2) wt_wait_cluster(&peer_full[s], ph, t0);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t sa = smem_u32(smem + s * STAGE);
                        const uint32_t a_lo = ((sa & 0x3FFFFu) >> 4) | ((128u >> 4) << 16);      // LBO 128 B
                        const uint32_t b_lo = (((sa + WH_TILE_BYTES) & 0x3FFFFu) >> 4) | ((128u >> 4) << 16);
#pragma unroll
                        for (int j = 0; j < 8; j++) {      // 8 x K=16: two core matrices (256 B) further along K each
                            if (a.dbg & 2) break;
                            uint64_t da, db;
                            asm("mov.b64 %0, {%1, %2};" : "=l"(da) : "r"(a_lo + j * 16), "r"(desc_hi));
                            asm("mov.b64 %0, {%1, %2};" : "=l"(db) : "r"(b_lo + j * 16), "r"(desc_hi));
                            wt_mma<CG>(tmem_base + u * WT_NT, da, db, idesc, (kt > 0 || j > 0) ? 1u : 0u);
                        }
                        wt_commit<CG>(&empty[s]);
                        if (kt == KT - 1) wt_commit<CG>(&acc_full[u]);
                    }
                    __syncwarp();
                }
            }
        } else {
            // second CTA of a pair: tell the leader when this CTA's half of a stage has landed
            const uint32_t leader_pf = mapa_shared(smem_u32(peer_full), 0);
            uint32_t it = 0;
            for (int t = cid; t < a.n_tiles; t += ncl) {
                for (int kt = 0; kt < KT; kt++, it++) {
                    const uint32_t s = it % NS;
                    wt_wait(&full[s], (it / NS) & 1, t0);
                    if (lane == 0) mbar_arrive_remote(leader_pf + s * 8);
                    __syncwarp();
                }
            }
        }
    } else {
        // ------------------------------------ epilogue ------------------------------------
        // warp = (TMEM lane quarter eq, group eg): group eg drains columns [8 eg, 8 eg + 8) of each 32-column chunk
        // into the smem tile and then handles activation rows eg, eg + 4, ... of the chunk
        const int eq = warp & 3, eg = (warp - 2) >> 2;
        const int etid = eq * 32 + lane;
        uint32_t tcount = 0;
        const uint32_t leader_ae = (CG == 2) ? mapa_shared(smem_u32(acc_empty), 0) : 0u;
        bool first = true;
        for (int t = cid; t < a.n_tiles; t += ncl, tcount++) {
            const int rp = t / m_tiles, mt = t - rp * m_tiles;
            const int rt = rp * CG + (int)crank;
            const int m0 = mt * WT_NT;
            const int Mt = min(WT_NT, a.M - m0);
            const uint32_t u = tcount & 1;
            wt_wait(&acc_full[u], (tcount >> 1) & 1, t0);
            tc_fence_after();
            if (first) { pdl_wait(); first = false; }
            const int n_chunks = (Mt + WT_EPI_CHUNK - 1) / WT_EPI_CHUNK;
            for (int ch = 0; ch < n_chunks; ch++) {
                const int mc = min(WT_EPI_CHUNK, Mt - ch * WT_EPI_CHUNK);
                const uint32_t taddr = tmem_base + ((uint32_t)(eq * 32) << 16) + u * WT_NT + (uint32_t)(ch * WT_EPI_CHUNK) + eg * 8;
                uint32_t v[8];
                tmem_ld_32x32b_x8(taddr, v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 8; j++) out_s[(eg * 8 + j) * 128 + etid] = __uint_as_float(v[j]);
                tc_fence_before();
                asm volatile("bar.sync 1, 512;");
                if (ch == n_chunks - 1 && tid == 64) {      // every warp has read its last columns: the accumulator is free
                    if (CG == 2 && crank != 0) mbar_arrive_remote(leader_ae + u * 8);
                    else mbar_arrive(&acc_empty[u]);
                }
                if (!(a.dbg & 1)) epilogue_rows(a.epi, out_s, 128, 128, rt * 128, a.d_real, mc, m0 + ch * WT_EPI_CHUNK, etid, eg, 4);
                asm volatile("bar.sync 1, 512;");
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) wt_cluster_sync();      // nobody leaves while the partner may still arrive on its barriers
    if (warp == 1) {
        tc_fence_after();
        wt_tmem_dealloc<CG>(tmem_base);
    }
    epilogue_complete(a.epi, tid, a.n_epi_ctas, sflag);
    trace_end(a.tr);
}

// ------------------------------------------------------------------------------------------------
// Q40T -> WH: one CTA per 128 x 128 tile, thread = (row, 32-block).  Same arithmetic as the dequantisers of
// gemm_q40_tc.cu: low nibbles ((1024 + e) - 1032) * d, high nibbles ((1024 + 16 e) - 1152) * (d / 16).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) q40t_to_wh_kernel(const uint8_t* __restrict__ q40t, uint8_t* __restrict__ wh) {
    const uint8_t* raw = q40t + (size_t)blockIdx.x * Q40T_TILE_BYTES;
    uint8_t* out = wh + (size_t)blockIdx.x * WH_TILE_BYTES;
    const int r = threadIdx.x & 127, kb = threadIdx.x >> 7;
    const uint4 nb = *reinterpret_cast<const uint4*>(raw + r * 64 + (((kb + (r >> 1)) & 3) << 4));
    const __half d = *reinterpret_cast<const __half*>(raw + Q40T_SCALE_OFF + r * 8 + kb * 2);
    const __half2 d2 = __half2half2(d);
    const __half2 d2h = __hmul2(d2, __float2half2_rn(0.0625f));
    const __half2 c1032 = __float2half2_rn(1032.0f), c1152 = __float2half2_rn(1152.0f);
    const uint32_t words[4] = {nb.x, nb.y, nb.z, nb.w};
    uint8_t* dst = out + (size_t)(r >> 3) * XH_ROWGROUP_BYTES + (size_t)(kb * 4) * 128 + (size_t)(r & 7) * 16;
#pragma unroll
    for (int w = 0; w < 4; w++) {
        const uint32_t x = words[w], x8 = x >> 8;
        uint32_t l02 = (x & 0x000F000Fu) | 0x64006400u, l13 = (x8 & 0x000F000Fu) | 0x64006400u;
        uint32_t h02 = (x & 0x00F000F0u) | 0x64006400u, h13 = (x8 & 0x00F000F0u) | 0x64006400u;
        __half2 v0 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&l02), c1032), d2);
        __half2 v1 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&l13), c1032), d2);
        __half2 v2 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&h02), c1152), d2h);
        __half2 v3 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&h13), c1152), d2h);
        uint4 o;
        o.x = *reinterpret_cast<uint32_t*>(&v0);
        o.y = *reinterpret_cast<uint32_t*>(&v1);
        o.z = *reinterpret_cast<uint32_t*>(&v2);
        o.w = *reinterpret_cast<uint32_t*>(&v3);
        *reinterpret_cast<uint4*>(dst + w * 128) = o;
    }
}

extern void dl_set_error(const char* fmt, ...);
extern "C" int dl_device_sm_count();

static int g_wh_dbg = 0, g_wh_pf_dist = 0, g_wh_pf_mask = 0, g_wh_piece = 16384;      // bytes per bulk copy: 16 KB pieces (4 / 8 KB measured: no gain)
extern "C" void dl_debug_set_wh_piece(int b) { if (b == 4096 || b == 8192 || b == 16384) g_wh_piece = b; }
extern "C" void dl_debug_set_wh_mode(int m) { g_wh_dbg = m; }
extern "C" void dl_debug_set_wh_prefetch(int dist, int mask) { g_wh_pf_dist = dist; g_wh_pf_mask = mask; }

extern "C" long long dl_wh_bytes(int d, int n) { return (long long)((d + 127) / 128) * (n / 128) * WH_TILE_BYTES; }

extern "C" int dl_q40t_to_wh(const uint8_t* q40t, int d, int n, uint8_t* wh, void* stream) {
    if (n % 128 != 0 || d < 1) { dl_set_error("dl_q40t_to_wh: n=%d must be a multiple of 128", n); return -1; }
    const int tiles = ((d + 127) / 128) * (n / 128);
    q40t_to_wh_kernel<<<tiles, 512, 0, (cudaStream_t)stream>>>(q40t, wh);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { dl_set_error("dl_q40t_to_wh: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}

// force_cg: 0 = pairs when the row-tile count is even, 1 = single CTAs (for tests and comparisons)
extern "C" int dl_gemm_wh_tc(const uint8_t* wh, int d, int n, const uint8_t* xh, int ncap, int M, const DlEpilogue* epi,
                             int force_cg, int pdl, void* stream) {
    if (n % 128 != 0) { dl_set_error("dl_gemm_wh_tc: n=%d not a multiple of 128", n); return -1; }
    if (M < 1 || ncap % 8 != 0 || ncap < M) { dl_set_error("dl_gemm_wh_tc: bad M=%d / ncap=%d", M, ncap); return -1; }
    if (epi->kind == DL_EPI_SILU_MUL && d % 64 != 0) { dl_set_error("dl_gemm_wh_tc: silu epilogue needs d%%64==0"); return -1; }
    WhArgs a;
    a.wh = wh; a.xh = xh; a.d_real = d;
    a.row_tiles = (d + 127) / 128;
    a.KT = n / 128; a.ncap = ncap; a.M = M;
    a.m_tiles = (M + WT_NT - 1) / WT_NT;
    const int cg = (force_cg == 1 || (a.row_tiles & 1)) ? 1 : 2;
    a.n_tiles = (a.row_tiles / cg) * a.m_tiles;
    const int n_sms = dl_device_sm_count();
    int grid = (n_sms / cg) * cg;
    if (grid > a.n_tiles * cg) grid = a.n_tiles * cg;
    a.n_epi_ctas = grid;
    a.epi = *epi;
    a.dbg = g_wh_dbg;
    a.pf_dist = g_wh_pf_dist < a.KT ? g_wh_pf_dist : 0;
    a.pf_mask = g_wh_pf_mask;
    a.piece = g_wh_piece;
    a.tr = dl_trace_next(TR_TC);
    const size_t smem = 196608 + WT_EPI_CHUNK * 128 * 4 + 16 * 8 + 16;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e1 = cudaFuncSetAttribute(gemm_wh_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaError_t e2 = cudaFuncSetAttribute(gemm_wh_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e1 != cudaSuccess || e2 != cudaSuccess) { dl_set_error("gemm wh smem attr: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : e2)); return -1; }
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(WT_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attrs[2];
    int na = 0;
    if (cg == 2) {
        attrs[na].id = cudaLaunchAttributeClusterDimension;
        attrs[na].val.clusterDim.x = 2;
        attrs[na].val.clusterDim.y = 1;
        attrs[na].val.clusterDim.z = 1;
        na++;
    }
    if (pdl) {
        attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attrs[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.attrs = attrs;
    cfg.numAttrs = na;
    cudaError_t e = (cg == 2) ? cudaLaunchKernelEx(&cfg, gemm_wh_tc_kernel<2>, a) : cudaLaunchKernelEx(&cfg, gemm_wh_tc_kernel<1>, a);
    if (e != cudaSuccess) { dl_set_error("gemm wh launch: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}
