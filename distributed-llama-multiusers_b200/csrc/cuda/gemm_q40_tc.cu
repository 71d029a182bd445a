// Q40 x Q80 -> f32 matmul on the 5th-gen tensor cores (tcgen05 / TMEM), M = 9 .. thousands of
// activation rows: batched multi-user decode and prompt prefill.
//
// The reference's exact arithmetic rescales every 32-element block by d_w*d_x (nn-cpu-ops.cpp:
// 352-377), which a TMEM accumulator cannot do between MMAs.  So the scales are folded into the
// operands: W tiles are dequantised Q40 -> fp16 ((nibble-8)*d, one rounding) inside the shared-memory
// pipeline, activations arrive as fp16(q*d) in the XH layout, and `tcgen05.mma.kind::f16`
// accumulates in f32 in TMEM.  "swap-AB": the 128 weight rows are the MMA M dimension (TMEM
// lanes), activation rows are the N dimension (16..128 per CTA), so small user batches still
// fill the tensor core's M.
//
// Warp roles (384 threads):
//   warp 0      producer: bulk async copies (UBLKCP) of raw 9 KB Q40T tiles and XH slabs
//   warp 1      TMEM allocator + single-thread MMA issuer (tcgen05.mma / tcgen05.commit)
//   warps 2-9   dequantisers: raw tile -> fp16 A tile in UMMA no-swizzle K-major layout
//   warps 2-5   (again) epilogue: tcgen05.ld TMEM -> registers -> smem tile -> fused epilogue
// Pipelines: rawA ring (producer -> dequant), B ring (producer -> MMA), A ring (dequant -> MMA,
// released by tcgen05.commit), accumulator barrier (MMA -> epilogue).  K can be split over a
// thread-block cluster (DSMEM reduction) when there are too few row tiles to fill the chip.
#include "epilogue.cuh"

#define TC_THREADS 384
#define TC_DEQ_WARPS 8
#define TC_A_BYTES 32768          // 128 rows x 128 K x fp16
#define TC_EPI_CHUNK 32           // activation rows per epilogue pass

struct TcArgs {
    const uint8_t* w;
    const uint8_t* xh;
    int d_real, d_pad, KT, ncap, M, NT, S;
    int NSA, NSB, NSH;
    int n_epi_ctas;
    unsigned long long* tr;
    DlEpilogue epi;
};

__global__ void __launch_bounds__(TC_THREADS, 1) gemm_q40_tc_kernel(const __grid_constant__ TcArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    trace_begin(a.tr);
    const int NT = a.NT;
    const int b_bytes = NT * 256;
    uint8_t* ringH = smem;                                          // [NSH][32 KB]
    uint8_t* ringB = ringH + a.NSH * TC_A_BYTES;                    // [NSB][NT*256]
    uint8_t* ringA = ringB + a.NSB * b_bytes;                       // [NSA][9216]
    float* out_s = reinterpret_cast<float*>(ringA + a.NSA * Q40T_TILE_BYTES);   // [32][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(out_s + TC_EPI_CHUNK * 128);
    uint64_t* fullA = bars;
    uint64_t* emptyA = fullA + 8;
    uint64_t* fullB = emptyA + 8;
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
    const uint32_t tmem_cols = NT < 32 ? 32u : (uint32_t)NT;

    if (tid == 0) {
        for (int s = 0; s < 8; s++) {
            mbar_init(&fullA[s], 1);
            mbar_init(&emptyA[s], TC_DEQ_WARPS);
            mbar_init(&fullB[s], 1);
            mbar_init(&emptyB[s], 1);
            mbar_init(&fullH[s], TC_DEQ_WARPS);
            mbar_init(&emptyH[s], 1);
        }
        mbar_init(accum_full, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_holder, tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
    pdl_launch_dependents();

    if (warp == 0) {
        // ------------------------------------ producer ------------------------------------
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            const uint8_t* wt = a.w + ((size_t)rt * a.KT + kt0) * Q40T_TILE_BYTES;
            const uint8_t* xb = a.xh + (size_t)kt0 * (a.ncap >> 3) * XH_ROWGROUP_BYTES + (size_t)(m0 >> 3) * XH_ROWGROUP_BYTES;
            const size_t x_kt_stride = (size_t)(a.ncap >> 3) * XH_ROWGROUP_BYTES;
            const int preA = nk < a.NSA ? nk : a.NSA;
            for (int i = 0; i < preA; i++) {      // weights first: independent of the previous kernel
                mbar_arrive_expect_tx(&fullA[i], Q40T_TILE_BYTES);
                bulk_g2s_hint(ringA + i * Q40T_TILE_BYTES, wt + (size_t)i * Q40T_TILE_BYTES, Q40T_TILE_BYTES, &fullA[i], pol);
            }
            pdl_wait();
            for (int i = 0; i < nk; i++) {
                const int sb = i % a.NSB;
                if (i >= a.NSB) mbar_wait(&emptyB[sb], ((i / a.NSB) & 1) ^ 1);
                mbar_arrive_expect_tx(&fullB[sb], b_bytes);
                bulk_g2s(ringB + sb * b_bytes, xb + (size_t)i * x_kt_stride, b_bytes, &fullB[sb]);
                const int ia = i + preA;
                if (ia < nk) {
                    const int sa = ia % a.NSA;
                    mbar_wait(&emptyA[sa], ((ia / a.NSA) & 1) ^ 1);
                    mbar_arrive_expect_tx(&fullA[sa], Q40T_TILE_BYTES);
                    bulk_g2s_hint(ringA + sa * Q40T_TILE_BYTES, wt + (size_t)ia * Q40T_TILE_BYTES, Q40T_TILE_BYTES, &fullA[sa], pol);
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------ MMA issuer ------------------------------------
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_f16_f32(128, NT);
            for (int i = 0; i < nk; i++) {
                const int sh = i % a.NSH, sb = i % a.NSB;
                mbar_wait(&fullH[sh], (i / a.NSH) & 1);
                mbar_wait(&fullB[sb], (i / a.NSB) & 1);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(ringH + sh * TC_A_BYTES);
                const uint32_t b_addr = smem_u32(ringB + sb * b_bytes);
#pragma unroll
                for (int j = 0; j < 8; j++) {      // 8 x K=16 per 128-wide k-tile; 2 core matrices (256 B) per step
                    const uint64_t da = umma_desc_kmajor_noswizzle(a_addr + j * 256, 128, 2048);
                    const uint64_t db = umma_desc_kmajor_noswizzle(b_addr + j * 256, 128, 2048);
                    umma_f16_ss(tmem_base, da, db, idesc, (i | j) != 0 ? 1u : 0u);
                }
                umma_commit(&emptyH[sh]);          // implicit tcgen05.fence::before_thread_sync
                umma_commit(&emptyB[sb]);
                if (i == nk - 1) umma_commit(accum_full);
            }
        }
        __syncwarp();
    } else if (warp < 2 + TC_DEQ_WARPS) {
        // ------------------------------------ dequantisers ------------------------------------
        const int t = tid - 64;                        // 0..255
        const __half2 c1032 = __float2half2_rn(1032.0f);
        const __half2 c16th = __float2half2_rn(0.0625f);
        const __half2 c72 = __float2half2_rn(72.0f);
        for (int i = 0; i < nk; i++) {
            const int sa = i % a.NSA, sh = i % a.NSH;
            mbar_wait(&fullA[sa], (i / a.NSA) & 1);
            if (i >= a.NSH) mbar_wait(&emptyH[sh], ((i / a.NSH) & 1) ^ 1);
            const uint8_t* raw = ringA + sa * Q40T_TILE_BYTES;
            uint8_t* At = ringH + sh * TC_A_BYTES;
#pragma unroll
            for (int half_i = 0; half_i < 2; half_i++) {
                const int it = t + half_i * 256;
                // a quarter-warp (8 lanes = one STS.128 / LDS.128 phase) covers 8 rows of ONE k-block: its
                // stores fill one contiguous 128-byte core matrix and its rotated raw loads hit 8 distinct
                // 16-byte slots -- both conflict-free (ncu: the (row%8,kb) lane order cost 4x the wavefronts)
                const int r = ((it >> 5) << 3) | (it & 7), kb = (it >> 3) & 3;
                const uint4 nb = *reinterpret_cast<const uint4*>(raw + r * 64 + (((kb + (r >> 1)) & 3) << 4));
                const __half dh = *reinterpret_cast<const __half*>(raw + Q40T_SCALE_OFF + r * 8 + kb * 2);
                const __half2 d2 = __half2half2(dh);
                uint8_t* dst = At + (r >> 3) * 2048 + (kb * 4) * 128 + (r & 7) * 16;
                const uint32_t words[4] = {nb.x, nb.y, nb.z, nb.w};
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    const uint32_t x = words[w];
                    const uint32_t x8 = x >> 8;
                    uint32_t l02 = (x & 0x000F000Fu) | 0x64006400u;      // half2(1024+e0, 1024+e2)
                    uint32_t l13 = (x8 & 0x000F000Fu) | 0x64006400u;     // half2(1024+e1, 1024+e3)
                    uint32_t h02 = (x & 0x00F000F0u) | 0x64006400u;      // half2(1024+16*e16, 1024+16*e18)
                    uint32_t h13 = (x8 & 0x00F000F0u) | 0x64006400u;
                    __half2 v0 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&l02), c1032), d2);
                    __half2 v1 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&l13), c1032), d2);
                    __half2 v2 = __hmul2(__hfma2(*reinterpret_cast<__half2*>(&h02), c16th, __hneg2(c72)), d2);
                    __half2 v3 = __hmul2(__hfma2(*reinterpret_cast<__half2*>(&h13), c16th, __hneg2(c72)), d2);
                    uint4 o;
                    o.x = *reinterpret_cast<uint32_t*>(&v0);
                    o.y = *reinterpret_cast<uint32_t*>(&v1);
                    o.z = *reinterpret_cast<uint32_t*>(&v2);
                    o.w = *reinterpret_cast<uint32_t*>(&v3);
                    *reinterpret_cast<uint4*>(dst + w * 128) = o;
                }
            }
            fence_proxy_async_smem();      // generic-proxy stores -> visible to the tensor core's async proxy
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&fullH[sh]);
                mbar_arrive(&emptyA[sa]);
            }
        }
    }

    // ---------------------------------------- epilogue ----------------------------------------
    // warps 2..5 own TMEM lane quarters (warp % 4); etid = output row of this thread
    const bool is_epi = (warp >= 2 && warp < 6);
    const int etid = (warp & 3) * 32 + lane;
    if (is_epi) {
        mbar_wait(accum_full, 0);
        tc_fence_after();
    }
    const int n_chunks = (Mt + TC_EPI_CHUNK - 1) / TC_EPI_CHUNK;
    for (int ch = 0; ch < n_chunks; ch++) {
        const int mc = min(TC_EPI_CHUNK, Mt - ch * TC_EPI_CHUNK);
        if (is_epi) {
            const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(ch * TC_EPI_CHUNK);
            const int ncol = min(TC_EPI_CHUNK, NT - ch * TC_EPI_CHUNK);
            for (int c0 = 0; c0 < ncol; c0 += 8) {
                uint32_t v[8];
                tmem_ld_32x32b_x8(taddr + c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 8; j++) out_s[(c0 + j) * 128 + etid] = __uint_as_float(v[j]);
            }
        }
        if (S > 1) {
            // split-K: only used with a single chunk (NT <= 32); all threads of the cluster sync
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
            if (crank == 0 && is_epi) pdl_wait();
            if (crank == 0 && is_epi) epilogue_rows(a.epi, out_s, 128, 128, rt * 128, a.d_real, mc, m0 + ch * TC_EPI_CHUNK, etid);
        } else {
            if (is_epi) {
                asm volatile("bar.sync 1, 128;");
                pdl_wait();
                epilogue_rows(a.epi, out_s, 128, 128, rt * 128, a.d_real, mc, m0 + ch * TC_EPI_CHUNK, etid);
                asm volatile("bar.sync 1, 128;");
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, tmem_cols);
    }
    if (crank == 0) epilogue_complete(a.epi, tid, a.n_epi_ctas, sflag);
    trace_end(a.tr);
}

extern void dl_set_error(const char* fmt, ...);

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
    if (ncap % a.NT != 0 && ncap < (M + a.NT - 1) / a.NT * a.NT) { dl_set_error("dl_gemm_q40_tc: ncap=%d too small for M=%d (tile %d)", ncap, M, a.NT); return -1; }
    if (a.NT > 32) k_split = 1;
    if (k_split != 1 && k_split != 2 && k_split != 4 && k_split != 8) { dl_set_error("dl_gemm_q40_tc: k_split must be 1,2,4,8"); return -1; }
    if (k_split > a.KT) k_split = 1;
    a.S = k_split;
    // stage counts by tile width: A(raw) 9 KB, B NT*256 B, H 32 KB, out 16 KB  (<= ~200 KB)
    if (a.NT <= 32) { a.NSA = 8; a.NSB = 4; a.NSH = 3; }
    else if (a.NT == 64) { a.NSA = 6; a.NSB = 3; a.NSH = 3; }
    else { a.NSA = 8; a.NSB = 2; a.NSH = 2; }
    a.n_epi_ctas = (a.d_pad / 128) * ((M + a.NT - 1) / a.NT);
    a.epi = *epi;
    a.tr = dl_trace_next(TR_TC);
    const size_t smem = (size_t)a.NSH * TC_A_BYTES + (size_t)a.NSB * a.NT * 256 + (size_t)a.NSA * Q40T_TILE_BYTES +
                        (size_t)TC_EPI_CHUNK * 128 * 4 + 6 * 8 * 8 + 8 + 16;
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
