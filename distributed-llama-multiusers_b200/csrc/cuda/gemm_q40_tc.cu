// Q40 x Q80 -> f32 matmul on the 5th-gen tensor cores (tcgen05 / TMEM), M = 9 .. thousands of
// activation rows: batched multi-user decode and prompt prefill.
//
// The reference's exact arithmetic rescales every 32-element block by d_w*d_x (nn-cpu-ops.cpp:
// 352-377), which a TMEM accumulator cannot do between MMAs.  So the scales are folded into the
// operands: W tiles are dequantised Q40 -> fp16 ((nibble-8)*d, one rounding) on the fly, activations
// arrive as fp16(q*d) in the XH layout, and `tcgen05.mma.kind::f16` accumulates in f32 in TMEM.
// "swap-AB": the 128 weight rows are the MMA M dimension (TMEM lanes), activation rows are the N
// dimension (16..256 per CTA), so small user batches still fill the tensor core's M.
//
// The dequantised weights never touch shared memory: the dequant warps read the raw 4.5-bit tile,
// expand it in registers and write it with tcgen05.st straight into TENSOR MEMORY, and the MMA takes
// its A operand from there (`tcgen05.mma [d], [a_tmem], b_desc`).  The first version staged fp16 A
// tiles in smem: 32 KB written + 32 KB read per 9 KB of weights saturated the 128 B/clk crossbar
// (per-tile timeline in profiles/: 1450 clk per k-tile, all of it in the MMA queue).
//
// Warp roles (672 threads):
//   warp 0       producer: bulk async copies (UBLKCP) of raw 9 KB Q40T tiles (lanes 0-3) and XH slabs (lanes 4-7)
//   warps 1,18-20 MMA issuers (one elected lane each; tcgen05.mma / tcgen05.commit), k-tiles round-robin,
//                one accumulator each; warp 1 also allocates TMEM
//   warps 2-17   dequantisers, 4 groups x 4 warps: group = k-tile % 4 = A stage, warp % 4 = TMEM lane
//                quarter, thread = weight row: LDS.128 x4 -> packed fp16x2 registers -> tcgen05.st.32x32b.x16
//   warps 2-17   (again) epilogue: tcgen05.ld TMEM -> registers -> smem tile -> fused epilogue (rows split 4 ways)
// TMEM columns: nacc f32 accumulators of NT columns from column 0, nstg 64-column A stages at the top
// (nacc*NT + 64*nstg <= 512: 4 acc + 6 stages for NT <= 32, 2 + 6 for NT = 64, 2 + 4 for 128, 1 + 4 for 256).
// Pipelines: raw ring (producer -> dequant), B ring (producer -> MMA), A stages (dequant -> MMA,
// released by tcgen05.commit), accumulator barrier (MMA -> epilogue).  K can be split over a
// thread-block cluster (DSMEM reduction) when there are too few row tiles to fill the chip.
#include "epilogue.cuh"

#define TC_DEQ_WARPS 16
#define TC_THREADS (64 + TC_DEQ_WARPS * 32 + 96)      // producer, MMA warp 0, dequantisers, MMA warps 1-3
#define TC_EPI_CHUNK 32           // activation rows per epilogue pass
#define TC_TMEM_COLS 512
#define TC_A_STAGE_COLS 64        // 128 K x fp16 = 64 32-bit columns per lane
#define TC_GROUPS 4               // dequantiser groups (k-tile i -> group i % 4)

struct TcArgs {
    const uint8_t* w;
    const uint8_t* xh;
    int d_real, d_pad, KT, ncap, M, NT, S;
    int NSA, NSB, nsb_log;
    int nstg;              // A stages in TMEM (64 columns each), at columns [512 - 64*nstg, 512)
    int nacc;              // MMA-issuing warps = independent TMEM accumulators (1, 2 or 4; nacc*NT <= 256), summed by the epilogue
    int LA, LB;            // issuing lanes per ring; each divides its stage count so a stage has ONE owner lane
    int n_epi_ctas;
    unsigned long long* tr;
    int dbg_mode;          // debug timing experiments (results invalid): 1 = one MMA per k-tile, 2 = no dequant/STTM
    long long* dbg;        // debug: per-tile clock64 stamps of CTA 0 (dl_debug_set_tc_trace), [tile][8]
    DlEpilogue epi;
};

#define TC_DBG(slot, i) do { if (a.dbg && blockIdx.x == 0 && blockIdx.y == 0 && (i) < 64) a.dbg[(i) * 8 + (slot)] = clock64(); } while (0)

__global__ void __launch_bounds__(TC_THREADS, 1) gemm_q40_tc_kernel(const __grid_constant__ TcArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    trace_begin(a.tr);
    if (threadIdx.x == 0) TC_DBG(0, 63);
    const int NT = a.NT;
    const int b_bytes = NT * 256;
    uint8_t* ringB = smem;                                          // [NSB][NT*256]
    uint8_t* ringA = ringB + a.NSB * b_bytes;                       // [NSA][9216]
    float* out_s = reinterpret_cast<float*>(ringA + a.NSA * Q40T_TILE_BYTES);   // [32][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(out_s + TC_EPI_CHUNK * 128);
    uint64_t* fullA = bars;                     // [16]
    uint64_t* emptyA = fullA + 16;              // [16]
    uint64_t* fullB = emptyA + 16;              // [8] each from here
    uint64_t* emptyB = fullB + 8;
    uint64_t* fullH = emptyB + 8;
    uint64_t* emptyH = fullH + 8;
    uint64_t* accum_full = emptyH + 8;
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(accum_full + 1);
    int* sflag = reinterpret_cast<int*>(tmem_holder + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int S = a.S;
    const int crank = (S > 1) ? (int)cg::this_cluster().block_rank() : 0;
    const int rt = blockIdx.x / S;
    const int m0 = blockIdx.y * NT;
    const int Mt = min(NT, a.M - m0);
    const int kt0 = (a.KT * crank) / S, kt1 = (a.KT * (crank + 1)) / S;
    const int nk = kt1 - kt0;
    const int nacc = (a.nacc <= nk) ? a.nacc : (nk >= 2 ? 2 : 1);      // accumulators in use = MMA-issuing warps (power of two)

    if (tid == 0) {
        for (int s = 0; s < 16; s++) {
            mbar_init(&fullA[s], 1);
            mbar_init(&emptyA[s], 4);
        }
        for (int s = 0; s < 8; s++) {
            mbar_init(&fullB[s], 1);
            mbar_init(&emptyB[s], 1);
            mbar_init(&fullH[s], 4);
            mbar_init(&emptyH[s], 1);
        }
        mbar_init(accum_full, nacc);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_holder, TC_TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
    if (threadIdx.x == 0) TC_DBG(1, 63);
    pdl_launch_dependents();

    if (warp == 0) {
        // ------------------------------------ producer ------------------------------------
        // lanes 0..LA-1 stream the raw weight tiles, lanes 4..4+LB-1 the activation slabs: one issuing thread
        // caps a CTA at ~34 GB/s (measured on the GEMV).  A stage has exactly one owner lane (LA | NSA,
        // LB | NSB): a stage shared by two lanes could be lapped by two phases, which a parity wait cannot see.
        // Every lane polls ITS next stage with a non-blocking test_wait inside one converged loop, so a lane
        // whose stage is still busy never holds back a lane whose stage is free (blocking waits serialise the
        // divergent lanes of a warp and let one role starve the other).
        const bool role_a = lane < a.LA, role_b = lane >= 4 && lane < 4 + a.LB;
        const uint8_t* wt = a.w + ((size_t)rt * a.KT + kt0) * Q40T_TILE_BYTES;
        const uint8_t* xb = a.xh + (size_t)kt0 * (a.ncap >> 3) * XH_ROWGROUP_BYTES + (size_t)(m0 >> 3) * XH_ROWGROUP_BYTES;
        const size_t x_kt_stride = (size_t)(a.ncap >> 3) * XH_ROWGROUP_BYTES;
        const uint64_t pol = l2_policy_evict_first();
        int i = role_a ? lane : (role_b ? lane - 4 : nk);
        const int step = role_a ? a.LA : a.LB;
        if (role_a) {
            // weights do not depend on the previous kernel: fill the whole ring before the PDL wait
            for (; i < nk && i < a.NSA; i += step) {
                if (!(a.dbg_mode & 4)) TC_DBG(0, i);
                mbar_arrive_expect_tx(&fullA[i], Q40T_TILE_BYTES);
                bulk_g2s_hint(ringA + i * Q40T_TILE_BYTES, wt + (size_t)i * Q40T_TILE_BYTES, Q40T_TILE_BYTES, &fullA[i], pol);
            }
        }
        __syncwarp();
        pdl_wait();
        while (__any_sync(0xffffffffu, i < nk)) {
            bool issued = false;
            if (i < nk) {
                if (role_a) {
                    const int sa = i % a.NSA;
                    if (mbar_test_wait(&emptyA[sa], ((i / a.NSA) & 1) ^ 1)) {      // i >= NSA here
                        if (!(a.dbg_mode & 4)) TC_DBG(0, i);
                        mbar_arrive_expect_tx(&fullA[sa], Q40T_TILE_BYTES);
                        bulk_g2s_hint(ringA + sa * Q40T_TILE_BYTES, wt + (size_t)i * Q40T_TILE_BYTES, Q40T_TILE_BYTES, &fullA[sa], pol);
                        i += step;
                        issued = true;
                    }
                } else {
                    const int sb = i % a.NSB;
                    if (i < a.NSB || mbar_test_wait(&emptyB[sb], ((i / a.NSB) & 1) ^ 1)) {
                        mbar_arrive_expect_tx(&fullB[sb], b_bytes);
                        bulk_g2s(ringB + sb * b_bytes, xb + (size_t)i * x_kt_stride, b_bytes, &fullB[sb]);
                        i += step;
                        issued = true;
                    }
                }
            }
            if (!__any_sync(0xffffffffu, issued)) __nanosleep(40);      // leave the issue slots to the dequantisers
        }
    } else if (warp == 1 || warp >= 2 + TC_DEQ_WARPS) {
        // ------------------------------------ MMA issuers ------------------------------------
        // One k-tile costs its issuing thread ~900 clk of mostly fixed latency (two barrier waits, fence,
        // 8 x ~45 clk UTCHMMA issue, two commits; per-tile timeline in profiles/), far more than the MMAs
        // themselves.  Up to four issuing warps take k-tiles round-robin, each into its OWN accumulator
        // (summed by the epilogue), so the result stays bit-deterministic.  Warp-uniform control flow (every
        // lane walks the barriers) with one elected lane issuing keeps the descriptors in uniform registers.
        const int mw = (warp == 1) ? 0 : warp - (2 + TC_DEQ_WARPS) + 1;
        const uint32_t idesc = umma_idesc_f16_f32(128, NT);
        // K-major no-swizzle descriptor of B: LBO 128 B (K core matrices), SBO 2048 B (8-row groups); only the
        // 14-bit start-address field changes between MMAs
        const uint32_t desc_hi = (2048u >> 4) | (1u << 14);
        const uint32_t desc_lo0 = ((smem_u32(ringB) & 0x3FFFFu) >> 4) | ((128u >> 4) << 16);
        const uint32_t acc = tmem_base + (uint32_t)(mw * NT);
        const bool det = (a.dbg_mode & 4) != 0;      // detail stamps of warp 1 only (slots 0..7)
        const bool stamp = (lane == 0 && mw == 0 && !(a.dbg_mode & 8));
        const int nsb_mask = a.NSB - 1, nsb_log = a.nsb_log;
        // narrow tiles are dequant-bound: the issuers sleep between probes and leave their sub-partition's issue
        // slots to the math warps; wide tiles are MMA-bound: the issuers are the critical path and spin
        const uint32_t mma_sleep = (NT <= 32) ? 64u : 0u;
        const int nstg = a.nstg;
        const uint32_t a_col0 = TC_TMEM_COLS - nstg * TC_A_STAGE_COLS;
        if (mw < nacc) {
            int sh = mw, ph = 0;      // A stage of k-tile i = i % nstg, phase parity = (i / nstg) & 1
            for (int i = mw; i < nk; i += nacc) {
                const int sb = i & nsb_mask, pb = (i >> nsb_log) & 1;
                if (det && stamp) TC_DBG(0, i);
                mbar_wait_sleep(&fullH[sh], ph, mma_sleep);
                if (stamp) TC_DBG(det ? 1 : 5, i);
                mbar_wait_sleep(&fullB[sb], pb, mma_sleep);
                if (stamp) TC_DBG(det ? 2 : 6, i);
                tc_fence_after();
                if (det && stamp) TC_DBG(3, i);
                if (elect_one()) {
                    const uint32_t a_tmem = tmem_base + a_col0 + sh * TC_A_STAGE_COLS;
                    const uint32_t lo = desc_lo0 + (uint32_t)((sb * b_bytes) >> 4);
#pragma unroll
                    for (int j = 0; j < 8; j++) {      // 8 x K=16 per 128-wide k-tile: 8 TMEM columns of A, 2 core matrices (256 B) of B
                        if ((a.dbg_mode & 1) && j > 0) break;
                        uint64_t db;
                        asm("mov.b64 %0, {%1, %2};" : "=l"(db) : "r"(lo + j * 16), "r"(desc_hi));
                        umma_f16_ts(acc, a_tmem + j * 8, db, idesc, (i != mw || j > 0) ? 1u : 0u);
                    }
                    if (det && mw == 0) TC_DBG(4, i);
                    umma_commit(&emptyH[sh]);          // implicit tcgen05.fence::before_thread_sync
                    umma_commit(&emptyB[sb]);
                    if (i + nacc >= nk) umma_commit(accum_full);      // this warp's last k-tile
                    if (det && mw == 0) TC_DBG(6, i);
                }
                __syncwarp();
                if (stamp) TC_DBG(7, i);
                sh += nacc;
                if (sh >= nstg) { sh -= nstg; ph ^= 1; }
            }
        }
    } else if (warp < 2 + TC_DEQ_WARPS) {
        // ------------------------------------ dequantisers ------------------------------------
        // Four groups of four warps; group g owns k-tiles g, g+4, ... and A stage g, so the groups' latency
        // chains (raw wait -> LDS -> unpack -> STTM -> wait::st -> arrive) overlap instead of all 16 warps
        // marching through every tile in lock-step.  Inside a group warp = TMEM lane quarter (warp % 4),
        // thread = weight row r, which it expands for the whole 128-wide k-tile (4 blocks of 32).
        const int q = warp & 3, g = (warp - 2) >> 2;
        const int r = q * 32 + lane;
        // consecutive rows rotate their four 16-byte k-block slots, so the 8 lanes of an LDS.128 phase hit
        // 8 distinct bank groups (same layout the GEMV reads row-per-lane)
        const uint32_t row_off = r * 64, rot = (uint32_t)(r >> 1);
        const uint32_t sc_off = Q40T_SCALE_OFF + r * 8;
        const int nstg = a.nstg;
        const uint32_t a_taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (TC_TMEM_COLS - nstg * TC_A_STAGE_COLS);
        const __half2 c1032 = __float2half2_rn(1032.0f);
        const __half2 c1152 = __float2half2_rn(1152.0f);
        const __half2 sixteenth = __float2half2_rn(0.0625f);
        const bool stamp = (tid == 64 && !(a.dbg_mode & 12));
        const bool qstamp = (lane == 0 && (a.dbg_mode & 8));      // mode 8: slots 1..4 = arrival time of lane quarter 0..3
        const int nsa = a.NSA;
        int sh = g, ph = 0;      // A stage of k-tile i = i % nstg (g < 4 <= nstg), phase parity = (i / nstg) & 1
        int sa = g, pa = 0;      // raw ring stage i % NSA and its phase parity
        for (int i = g; i < nk; i += TC_GROUPS) {
            mbar_wait_sleep(&fullA[sa], pa, 32);
            if (stamp) TC_DBG(1, i);
            const uint8_t* raw = ringA + sa * Q40T_TILE_BYTES;
            uint4 nb[4];
#pragma unroll
            for (int kb = 0; kb < 4; kb++) nb[kb] = *reinterpret_cast<const uint4*>(raw + row_off + (((kb + rot) & 3) << 4));
            const uint2 sc = *reinterpret_cast<const uint2*>(raw + sc_off);      // 4 fp16 scales of this row
            // the tile now lives in registers: hand the raw stage back BEFORE the unpack (the ring has to cover
            // DRAM latency + queueing; holding a stage through the ~1300 clk unpack halved its effective depth).
            // The loads above were issued by this warp before the arrive and shared-memory operations of a warp
            // complete in order; the refill is >500 ns away (barrier -> producer poll -> HBM fetch).
            __syncwarp();
            if (lane == 0) mbar_arrive(&emptyA[sa]);
            if (i >= nstg) mbar_wait_sleep(&emptyH[sh], ph ^ 1, 32);      // the MMAs that read this A stage have retired
            if (stamp) TC_DBG(2, i);
            tc_fence_after();
#pragma unroll
            for (int kb = 0; kb < 4; kb++) {
                const uint32_t sw = (kb < 2) ? sc.x : sc.y;
                const uint32_t s16 = (kb & 1) ? (sw >> 16) : (sw & 0xFFFFu);
                const uint32_t sdup = s16 | (s16 << 16);
                const __half2 d2 = *reinterpret_cast<const __half2*>(&sdup);
                const __half2 d2h = __hmul2(d2, sixteenth);      // exact (power of two) unless d is subnormal
                const uint32_t words[4] = {nb[kb].x, nb[kb].y, nb[kb].z, nb[kb].w};
                uint32_t v[16];
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    // K-slot order inside the 32-block (XH uses the same): e(4w), e(4w+2), e(4w+1), e(4w+3), then +16
                    const uint32_t x = words[w];
                    const uint32_t x8 = x >> 8;
                    uint32_t l02 = (x & 0x000F000Fu) | 0x64006400u;      // half2(1024+e0, 1024+e2)
                    uint32_t l13 = (x8 & 0x000F000Fu) | 0x64006400u;     // half2(1024+e1, 1024+e3)
                    uint32_t h02 = (x & 0x00F000F0u) | 0x64006400u;      // half2(1024+16*e16, 1024+16*e18)
                    uint32_t h13 = (x8 & 0x00F000F0u) | 0x64006400u;
                    __half2 v0 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&l02), c1032), d2);
                    __half2 v1 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&l13), c1032), d2);
                    __half2 v2 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&h02), c1152), d2h);   // 16(e-8) * d/16
                    __half2 v3 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&h13), c1152), d2h);
                    v[w * 4 + 0] = *reinterpret_cast<uint32_t*>(&v0);
                    v[w * 4 + 1] = *reinterpret_cast<uint32_t*>(&v1);
                    v[w * 4 + 2] = *reinterpret_cast<uint32_t*>(&v2);
                    v[w * 4 + 3] = *reinterpret_cast<uint32_t*>(&v3);
                }
                if (!(a.dbg_mode & 2)) tmem_st_32x32b_x16(a_taddr + sh * TC_A_STAGE_COLS + kb * 16, v);
            }
            tmem_st_wait();
            if (stamp) TC_DBG(3, i);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&fullH[sh]);
            if (stamp) TC_DBG(4, i);
            if (qstamp) TC_DBG(1 + q, i);
            sh += TC_GROUPS;
            if (sh >= nstg) { sh -= nstg; ph ^= 1; }
            sa += TC_GROUPS;
            if (sa >= nsa) { sa -= nsa; pa ^= 1; }
        }
    }

    // ---------------------------------------- epilogue ----------------------------------------
    // all 16 dequant warps: warp = (TMEM lane quarter eq, group eg); group eg drains columns [8eg, 8eg+8) of
    // each 32-column chunk into the smem tile, then handles activation rows eg, eg+4, ... of the chunk
    const bool is_epi = (warp >= 2 && warp < 2 + TC_DEQ_WARPS);
    const int eq = warp & 3, eg = (warp - 2) >> 2;
    const int etid = eq * 32 + lane;      // output row of this thread
    if (is_epi) {
        mbar_wait(accum_full, 0);
        tc_fence_after();
        if (tid == 64) TC_DBG(2, 63);
    }
    const int n_chunks = (Mt + TC_EPI_CHUNK - 1) / TC_EPI_CHUNK;
    for (int ch = 0; ch < n_chunks; ch++) {
        const int mc = min(TC_EPI_CHUNK, Mt - ch * TC_EPI_CHUNK);
        if (is_epi) {
            const uint32_t taddr = tmem_base + ((uint32_t)(eq * 32) << 16) + (uint32_t)(ch * TC_EPI_CHUNK);
            const int ncol = min(TC_EPI_CHUNK, NT - ch * TC_EPI_CHUNK);
            const int c0 = eg * 8;
            if (c0 < ncol) {
                uint32_t v[8];
                tmem_ld_32x32b_x8(taddr + c0, v);
                tmem_ld_wait();
                float f[8];
#pragma unroll
                for (int j = 0; j < 8; j++) f[j] = __uint_as_float(v[j]);
                for (int ac = 1; ac < nacc; ac++) {
                    tmem_ld_32x32b_x8(taddr + c0 + ac * NT, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 8; j++) f[j] += __uint_as_float(v[j]);
                }
#pragma unroll
                for (int j = 0; j < 8; j++) out_s[(c0 + j) * 128 + etid] = f[j];
            }
        }
        if (S > 1) {
            // split-K: every rank publishes its partial chunk, the leader sums them over DSMEM and runs the
            // epilogue; all threads of the cluster take part in both syncs
            cg::cluster_group cluster = cg::this_cluster();
            tc_fence_before();
            cluster.sync();
            if (crank == 0) {
                for (int idx = tid; idx < mc * 128; idx += TC_THREADS) {
                    float v = out_s[idx];
                    for (int c = 1; c < S; c++) v += cluster.map_shared_rank(out_s, c)[idx];
                    out_s[idx] = v;
                }
            }
            cluster.sync();
            if (crank == 0 && is_epi) {
                pdl_wait();
                epilogue_rows(a.epi, out_s, 128, 128, rt * 128, a.d_real, mc, m0 + ch * TC_EPI_CHUNK, etid, eg, TC_GROUPS);
                if (ch + 1 < n_chunks) asm volatile("bar.sync 1, 512;");      // out_s is rewritten by the next chunk
            }
        } else if (is_epi) {
            if (tid == 64 && ch == 0) TC_DBG(4, 63);
            asm volatile("bar.sync 1, 512;");
            pdl_wait();
            if (tid == 64 && ch == 0) TC_DBG(5, 63);
            epilogue_rows(a.epi, out_s, 128, 128, rt * 128, a.d_real, mc, m0 + ch * TC_EPI_CHUNK, etid, eg, TC_GROUPS);
            if (tid == 64 && ch == 0) TC_DBG(6, 63);
            asm volatile("bar.sync 1, 512;");
            if (tid == 64 && ch == 0) TC_DBG(7, 63);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TC_TMEM_COLS);
    }
    if (crank == 0) epilogue_complete(a.epi, tid, a.n_epi_ctas, sflag);
    if (threadIdx.x == 0) TC_DBG(3, 63);
    trace_end(a.tr);
}

extern void dl_set_error(const char* fmt, ...);
extern "C" int dl_device_sm_count();

static long long* g_tc_dbg = nullptr;
static int g_tc_dbg_mode = 0;
extern "C" void dl_debug_set_tc_mode(int m) { g_tc_dbg_mode = m; }
extern "C" void dl_debug_set_tc_trace(long long* p) { g_tc_dbg = p; }

extern "C" int dl_gemm_q40_tc(const uint8_t* w_q40t, int d, int n, const uint8_t* xh, int ncap, int M, int k_split,
                              const DlEpilogue* epi, int pdl, void* stream) {
    if (n % Q40T_TILE_K != 0) { dl_set_error("dl_gemm_q40_tc: n=%d not a multiple of 128", n); return -1; }
    if (M < 1) { dl_set_error("dl_gemm_q40_tc: M < 1"); return -1; }
    if (epi->kind == DL_EPI_SILU_MUL && d % 64 != 0) { dl_set_error("dl_gemm_q40_tc: silu epilogue needs d%%64==0"); return -1; }
    TcArgs a;
    a.w = w_q40t; a.xh = xh;
    a.d_real = d;
    a.d_pad = (d + 127) / 128 * 128;
    a.KT = n / Q40T_TILE_K;
    a.ncap = ncap; a.M = M;
    a.NT = (M <= 16) ? 16 : (M <= 32) ? 32 : (M <= 64) ? 64 : 128;
    const int row_tiles = (d + 127) / 128, n_sms = dl_device_sm_count();
    const int ks = (k_split == 2 || k_split == 4 || k_split == 8) ? k_split : 1;
    // 256-wide tiles halve the dequant work per activation row, but only pay when they still leave enough CTAs
    // to fill the chip (w13: 224 row tiles) ...
    if (M > 128 && ncap % 256 == 0 && row_tiles * ((M + 255) / 256) * ks >= n_sms / 2) a.NT = 256;
    // ... and narrow matrices (qkv / wo: 32-48 row tiles) are better off with narrower tiles that launch more
    // CTAs, even though every CTA then unpacks the weight tile again (the unpack overlaps the MMAs)
    while (a.NT > 32 && row_tiles * ((M + a.NT - 1) / a.NT) * ks < n_sms / 2) a.NT >>= 1;
    if (ncap < (M + a.NT - 1) / a.NT * a.NT) { dl_set_error("dl_gemm_q40_tc: ncap=%d too small for M=%d (tile %d)", ncap, M, a.NT); return -1; }
    if (k_split != 1 && k_split != 2 && k_split != 4 && k_split != 8) { dl_set_error("dl_gemm_q40_tc: k_split must be 1,2,4,8"); return -1; }
    if (k_split > a.KT) k_split = 1;
    a.S = k_split;
    // stage counts by tile width: raw 9 KB, B NT*256 B, out 16 KB  (<= ~216 KB)
    if (a.NT <= 64) { a.NSA = 12; a.NSB = 4; }
    else if (a.NT == 128) { a.NSA = 8; a.NSB = 4; }
    else { a.NSA = 8; a.NSB = 2; }
    if (a.NSA < TC_GROUPS || a.NSA > 16) { dl_set_error("dl_gemm_q40_tc: bad NSA=%d", a.NSA); return -1; }
    // nacc must divide nstg: then successive phases of an A stage are consumed by the SAME issuing warp (and
    // written after a retire that is ordered behind the previous one), so a parity wait can never be lapped
    a.nacc = (a.NT <= 128) ? 2 : 1;
    a.nstg = (a.NT <= 64) ? 6 : 4;
    if (a.nacc * a.NT + a.nstg * TC_A_STAGE_COLS > TC_TMEM_COLS || a.nstg < TC_GROUPS || a.nstg > 8 || a.nstg % a.nacc != 0 || (a.nacc != 1 && a.nacc != 2 && a.nacc != 4)) {
        dl_set_error("dl_gemm_q40_tc: bad TMEM plan nacc=%d nstg=%d NT=%d", a.nacc, a.nstg, a.NT);
        return -1;
    }
    a.nsb_log = (a.NSB == 8) ? 3 : (a.NSB == 4) ? 2 : 1;
    a.LA = (a.NSA % 4 == 0) ? 4 : (a.NSA % 3 == 0) ? 3 : (a.NSA % 2 == 0) ? 2 : 1;
    a.LB = (a.NSB % 4 == 0) ? 4 : (a.NSB % 3 == 0) ? 3 : (a.NSB % 2 == 0) ? 2 : 1;
    a.n_epi_ctas = (a.d_pad / 128) * ((M + a.NT - 1) / a.NT);
    a.epi = *epi;
    a.tr = dl_trace_next(TR_TC);
    a.dbg = g_tc_dbg;
    a.dbg_mode = g_tc_dbg_mode;
    const size_t smem = (size_t)a.NSB * a.NT * 256 + (size_t)a.NSA * Q40T_TILE_BYTES +
                        (size_t)TC_EPI_CHUNK * 128 * 4 + (2 * 16 + 4 * 8) * 8 + 8 + 16;
    static size_t max_set = 0;
    if (smem > max_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_q40_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { dl_set_error("gemm tc smem attr: %s", cudaGetErrorString(e)); return -(int)e; }
        max_set = smem;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((a.d_pad / 128) * a.S), (unsigned)((M + a.NT - 1) / a.NT));
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attrs[2];
    int na = 0;
    if (a.S > 1) {
        attrs[na].id = cudaLaunchAttributeClusterDimension;
        attrs[na].val.clusterDim.x = a.S;
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
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_q40_tc_kernel, a);
    if (e != cudaSuccess) { dl_set_error("gemm tc launch: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}
