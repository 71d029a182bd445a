// Q40 x Q80 -> f32 matmul for small activation batches (M <= 8): the decode hot path.
//
// Design (B200-first, not a port of matmul_Q80_Q40_F32, src/nn/nn-cpu-ops.cpp:222-440):
//  * weights live in HBM as Q40T tiles (common.cuh): each pipeline stage is ONE contiguous
//    bulk async copy (cp.async.bulk -> SASS UBLKCP) of a 128x128 tile part into shared memory,
//    completion tracked by mbarriers; an 8-stage ring keeps ~80 KB per SM in flight, which is
//    what it takes to cover HBM3e latency at 6.5 TB/s over 148 SMs.
//  * exact reference arithmetic: int8 dot products (dp4a) per 32-block, one f32 multiply by
//    d_w*d_x, f32 accumulation.
//  * grid = row_tiles * R (row split) * S (K split); the S CTAs of a tile form a thread-block
//    cluster and reduce their partial sums through distributed shared memory -- deterministic,
//    no atomics, no global scratch.
//  * weight prefetch is issued before griddepcontrol.wait, so under programmatic dependent
//    launch the first stages stream while the previous kernel (or a peer GPU) still runs.
//  * the epilogue (epilogue.cuh) fuses RoPE + KV-cache write, SiLU*mul + Q80 cast, or the
//    Q80-cast partial-sum push to every peer GPU with release flags.
#include "epilogue.cuh"

#define GEMV_NS 8                       // pipeline stages
#define GEMV_CONSUMERS 256              // 2 groups x 128 threads
#define GEMV_THREADS (GEMV_CONSUMERS + 32)
#define GEMV_MAX_M 8
#define GEMV_STAGE_BYTES (Q40T_TILE_BYTES + GEMV_MAX_M * XQ_ROW_BYTES)   // 10496

struct GemvArgs {
    const uint8_t* w;
    const uint8_t* xq;
    int d_real, d_pad, KT, mcap, M, R, S;
    int n_epi_ctas;
    DlEpilogue epi;
};

template <int MT>
__global__ void __launch_bounds__(GEMV_THREADS, 1) gemv_q40_kernel(const __grid_constant__ GemvArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* stages = smem;
    float* red = reinterpret_cast<float*>(smem + GEMV_NS * GEMV_STAGE_BYTES);     // [2*R][MT][rows]  (<= 2*MT*128)
    float* out_s = red + 2 * MT * 128;                                            // [MT][128]
    uint64_t* full = reinterpret_cast<uint64_t*>(out_s + MT * 128);
    uint64_t* empty = full + GEMV_NS;
    int* sflag = reinterpret_cast<int*>(empty + GEMV_NS);

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int S = a.S, R = a.R;
    const int crank = (S > 1) ? (int)cg::this_cluster().block_rank() : 0;
    const int tile_idx = blockIdx.x / S;
    const int rt = tile_idx / R, rs = tile_idx % R;
    const int rows = Q40T_TILE_ROWS / R;
    const int kt0 = (a.KT * crank) / S, kt1 = (a.KT * (crank + 1)) / S;
    const int nk = kt1 - kt0;

    if (tid == 0) {
        for (int s = 0; s < GEMV_NS; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 4);
        }
        mbar_fence_init();
    }
    __syncthreads();
    pdl_launch_dependents();

    const uint32_t nib_bytes = rows * 64, sc_bytes = rows * 8, act_bytes = a.M * XQ_ROW_BYTES;

    if (warp == GEMV_CONSUMERS / 32) {
        // ------------------------------ producer warp ------------------------------
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            const uint8_t* wt = a.w + ((size_t)rt * a.KT + kt0) * Q40T_TILE_BYTES;
            const int pre = nk < GEMV_NS ? nk : GEMV_NS;
            // weights do not depend on the previous kernel: stream them before the PDL wait
            for (int i = 0; i < pre; i++) {
                uint8_t* st = stages + i * GEMV_STAGE_BYTES;
                const uint8_t* t = wt + (size_t)i * Q40T_TILE_BYTES;
                mbar_arrive_expect_tx(&full[i], nib_bytes + sc_bytes + act_bytes);
                bulk_g2s_hint(st, t + rs * nib_bytes, nib_bytes, &full[i], pol);
                bulk_g2s_hint(st + Q40T_NIB_BYTES, t + Q40T_SCALE_OFF + rs * sc_bytes, sc_bytes, &full[i], pol);
            }
            pdl_wait();
            for (int i = 0; i < pre; i++) {
                uint8_t* st = stages + i * GEMV_STAGE_BYTES;
                bulk_g2s(st + Q40T_TILE_BYTES, a.xq + (size_t)(kt0 + i) * a.mcap * XQ_ROW_BYTES, act_bytes, &full[i]);
            }
            for (int i = pre; i < nk; i++) {
                const int s = i % GEMV_NS;
                mbar_wait(&empty[s], ((i / GEMV_NS) & 1) ^ 1);
                uint8_t* st = stages + s * GEMV_STAGE_BYTES;
                const uint8_t* t = wt + (size_t)i * Q40T_TILE_BYTES;
                mbar_arrive_expect_tx(&full[s], nib_bytes + sc_bytes + act_bytes);
                bulk_g2s_hint(st, t + rs * nib_bytes, nib_bytes, &full[s], pol);
                bulk_g2s_hint(st + Q40T_NIB_BYTES, t + Q40T_SCALE_OFF + rs * sc_bytes, sc_bytes, &full[s], pol);
                bulk_g2s(st + Q40T_TILE_BYTES, a.xq + (size_t)(kt0 + i) * a.mcap * XQ_ROW_BYTES, act_bytes, &full[s]);
            }
        }
    } else {
        // ------------------------------ consumers ------------------------------
        const int g = tid >> 7;              // consumer group: k-iterations i with i%2 == g
        const int t = tid & 127;
        const int row = t % rows;            // local row
        const int kq = t / rows;             // handles blocks kb = kq, kq+R, ...
        const int rtile = rs * rows + row;   // row inside the 128-row tile (for the rotation)
        float acc[MT];
#pragma unroll
        for (int m = 0; m < MT; m++) acc[m] = 0.0f;

        for (int i = g; i < nk; i += 2) {
            const int s = i % GEMV_NS;
            mbar_wait(&full[s], (i / GEMV_NS) & 1);
            const uint8_t* st = stages + s * GEMV_STAGE_BYTES;
            const uint8_t* act = st + Q40T_TILE_BYTES;
            for (int kb = kq; kb < 4; kb += R) {
                const uint4 nb = *reinterpret_cast<const uint4*>(st + row * 64 + (((kb + (rtile >> 1)) & 3) << 4));
                const float dw = __half2float(*reinterpret_cast<const __half*>(st + Q40T_NIB_BYTES + row * 8 + kb * 2));
                const int lo0 = nb.x & 0x0F0F0F0F, hi0 = (nb.x >> 4) & 0x0F0F0F0F;
                const int lo1 = nb.y & 0x0F0F0F0F, hi1 = (nb.y >> 4) & 0x0F0F0F0F;
                const int lo2 = nb.z & 0x0F0F0F0F, hi2 = (nb.z >> 4) & 0x0F0F0F0F;
                const int lo3 = nb.w & 0x0F0F0F0F, hi3 = (nb.w >> 4) & 0x0F0F0F0F;
#pragma unroll
                for (int m = 0; m < MT; m++) {
                    if (m < a.M) {
                        const uint8_t* xr = act + m * XQ_ROW_BYTES;
                        const int4 xl = *reinterpret_cast<const int4*>(xr + kb * 32);
                        const int4 xh = *reinterpret_cast<const int4*>(xr + kb * 32 + 16);
                        int isum = __dp4a(lo0, xl.x, 0);
                        isum = __dp4a(lo1, xl.y, isum);
                        isum = __dp4a(lo2, xl.z, isum);
                        isum = __dp4a(lo3, xl.w, isum);
                        isum = __dp4a(hi0, xh.x, isum);
                        isum = __dp4a(hi1, xh.y, isum);
                        isum = __dp4a(hi2, xh.z, isum);
                        isum = __dp4a(hi3, xh.w, isum);
                        const float dx = reinterpret_cast<const float*>(xr + 128)[kb];
                        const int s8 = reinterpret_cast<const int*>(xr + 144)[kb];
                        acc[m] += (float)(isum - s8) * (dw * dx);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
        // partial sums of the 2*R (group, kq) owners of each row
#pragma unroll
        for (int m = 0; m < MT; m++) red[((g * R + kq) * MT + m) * rows + row] = acc[m];
    }
    __syncthreads();
    for (int idx = tid; idx < MT * rows; idx += GEMV_THREADS) {
        const int m = idx / rows, r = idx % rows;
        float v = 0.0f;
        for (int j = 0; j < 2 * R; j++) v += red[(j * MT + m) * rows + r];
        out_s[m * 128 + r] = v;
    }
    if (S > 1) {
        cg::cluster_group cluster = cg::this_cluster();
        cluster.sync();
        if (crank == 0) {
            for (int idx = tid; idx < MT * rows; idx += GEMV_THREADS) {
                const int m = idx / rows, r = idx % rows;
                float v = out_s[m * 128 + r];
                for (int c = 1; c < S; c++) v += cluster.map_shared_rank(out_s, c)[m * 128 + r];
                out_s[m * 128 + r] = v;
            }
        }
        cluster.sync();
        if (crank != 0) return;
    }
    __syncthreads();
    pdl_wait();
    epilogue_rows(a.epi, out_s, 128, rows, rt * Q40T_TILE_ROWS + rs * rows, a.d_real, a.M, 0, tid);
    epilogue_complete(a.epi, tid, a.n_epi_ctas, sflag);
}

static size_t gemv_smem_bytes(int MT) {
    return (size_t)GEMV_NS * GEMV_STAGE_BYTES + (size_t)(2 * MT * 128 + MT * 128) * 4 + 2 * GEMV_NS * 8 + 16;
}

extern void dl_set_error(const char* fmt, ...);

template <int MT>
static int launch_gemv(const GemvArgs& a, int pdl, cudaStream_t stream) {
    static bool attr_set = false;
    const size_t smem = gemv_smem_bytes(MT);
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemv_q40_kernel<MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { dl_set_error("gemv smem attr: %s", cudaGetErrorString(e)); return -(int)e; }
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((a.d_pad / Q40T_TILE_ROWS) * a.R * a.S));
    cfg.blockDim = dim3(GEMV_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
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
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemv_q40_kernel<MT>, a);
    if (e != cudaSuccess) { dl_set_error("gemv launch: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}

extern "C" int dl_gemv_q40(const uint8_t* w_q40t, int d, int n, const uint8_t* xq, int mcap, int M, int row_split,
                           int k_split, const DlEpilogue* epi, int pdl, void* stream) {
    if (M < 1 || M > GEMV_MAX_M) { dl_set_error("dl_gemv_q40: M=%d out of range 1..8", M); return -1; }
    if (n % Q40T_TILE_K != 0) { dl_set_error("dl_gemv_q40: n=%d not a multiple of 128", n); return -1; }
    if (row_split != 1 && row_split != 2 && row_split != 4) { dl_set_error("dl_gemv_q40: row_split must be 1,2,4"); return -1; }
    if (k_split != 1 && k_split != 2 && k_split != 4 && k_split != 8) { dl_set_error("dl_gemv_q40: k_split must be 1,2,4,8"); return -1; }
    if (epi->kind == DL_EPI_SILU_MUL && (row_split > 2 || d % 64 != 0)) { dl_set_error("dl_gemv_q40: silu epilogue needs row_split<=2, d%%64==0"); return -1; }
    GemvArgs a;
    a.w = w_q40t;
    a.xq = xq;
    a.d_real = d;
    a.d_pad = (d + Q40T_TILE_ROWS - 1) / Q40T_TILE_ROWS * Q40T_TILE_ROWS;
    a.KT = n / Q40T_TILE_K;
    if (k_split > a.KT) k_split = 1;
    a.mcap = mcap;
    a.M = M;
    a.R = row_split;
    a.S = k_split;
    a.n_epi_ctas = (a.d_pad / Q40T_TILE_ROWS) * a.R;
    a.epi = *epi;
    cudaStream_t st = (cudaStream_t)stream;
    if (M == 1) return launch_gemv<1>(a, pdl, st);
    if (M == 2) return launch_gemv<2>(a, pdl, st);
    if (M <= 4) return launch_gemv<4>(a, pdl, st);
    return launch_gemv<8>(a, pdl, st);
}
