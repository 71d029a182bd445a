// Q40 x Q80 -> f32 matmul for small activation batches (M <= 8): the decode hot path.
//
// Design (B200-first, not a port of matmul_Q80_Q40_F32, src/nn/nn-cpu-ops.cpp:222-440):
//  * weights live in HBM as Q40T tiles (common.cuh): each pipeline stage is a contiguous bulk
//    async copy (cp.async.bulk -> SASS UBLKCP) of a 128x128 tile into shared memory, completion
//    tracked by mbarriers; a 12-stage ring keeps ~110 KB per SM in flight (16 MB chip-wide), which
//    covers HBM3e latency at 6.5 TB/s.
//  * one CTA per SM, 16 consumer warps in 4 groups (k-tiles are dealt round-robin to the groups)
//    plus a producer warp: the dp4a loop is latency-bound per warp, 4 warps per scheduler hide it.
//  * exact reference arithmetic: int8 dot products (dp4a) per 32-block, one f32 multiply by
//    d_w*d_x, f32 accumulation.
//  * the Q80 activations of this CTA's K range are resident in shared memory.  They either arrive
//    with one bulk copy (XQ slabs written by the previous kernel's epilogue), or -- fused prologue
//    -- the kernel builds them itself: wait for the peers' release flags, merge-add the partial
//    sums of all ranks into the residual stream, RMSNorm, Q80 cast (DlNormProlog).  The weight
//    prefetch is already in flight while this runs, so the all-gather / merge-add / norm / cast
//    of the reference (4 ops + a socket sync) costs no extra kernel and no global round trip.
//  * work split: a CTA owns 1..T row tiles (T > 1 for matrices with more row tiles than SMs: the
//    pipeline streams straight through tile boundaries) or, for small matrices, 1/S of the K range
//    of one row tile; the S CTAs of a tile form a thread-block cluster that shares the row's sum
//    of squares and pushes partial dot products into the leader through distributed shared
//    memory (remote stores + a remote mbarrier arrive) -- deterministic, no atomics, no global
//    scratch, no second cluster barrier.
//  * weight prefetch is issued before griddepcontrol.wait, so under programmatic dependent
//    launch the first stages stream while the previous kernel (or a peer GPU) still runs.
//  * the epilogue (epilogue.cuh) fuses RoPE + KV-cache write, SiLU*mul + Q80 cast, or the
//    Q80-cast partial-sum push to every peer GPU with release flags.
#include "epilogue.cuh"

#define GEMV_NS 12                      // pipeline stages
#define GEMV_NG 4                       // consumer groups of 128 threads
#define GEMV_NPROD 4                    // issuing lanes of the producer warp
#define GEMV_CONSUMERS (GEMV_NG * 128)
#define GEMV_THREADS (GEMV_CONSUMERS + 32)
#define GEMV_WARPS (GEMV_THREADS / 32)
#define GEMV_MAX_M 8

#define GEMV_TRACE(k)                                                                                      \
    do {                                                                                                   \
        if (a.trace != nullptr && (threadIdx.x == 0)) a.trace[blockIdx.x * 8 + (k)] = globaltimer_ns();    \
    } while (0)

struct GemvArgs {
    const uint8_t* w;
    const uint8_t* xq;
    int d_real, d_pad, KT, M, R, S, T, n;
    int n_units;              // row_tiles * R
    int n_epi_ctas;
    int fused;
    unsigned long long* trace;
    unsigned long long* tr;
    DlNormProlog norm;
    DlEpilogue epi;
};

DL_DEVINL void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(GEMV_CONSUMERS) : "memory"); }

// Q80-quantise one 32-element block held in v[8] (already normalised) and store it into the
// resident activation slab of k-block b, row m.
DL_DEVINL void store_q80_block_smem(const float4 (&v)[8], uint8_t* act_s, int M, int m, int b) {
    float amax = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; j++)
        amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v[j].x), fabsf(v[j].y)), fmaxf(fabsf(v[j].z), fabsf(v[j].w))));
    const float d = amax / 127.0f;
    const float id = (d != 0.0f) ? (1.0f / d) : 0.0f;
    int sumq = 0;
    uint32_t packed[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int q0 = __float2int_rn(v[j].x * id), q1 = __float2int_rn(v[j].y * id);
        const int q2 = __float2int_rn(v[j].z * id), q3 = __float2int_rn(v[j].w * id);
        sumq += q0 + q1 + q2 + q3;
        packed[j] = (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8) | ((uint32_t)(q2 & 0xff) << 16) |
                    ((uint32_t)(q3 & 0xff) << 24);
    }
    uint8_t* rowp = act_s + ((size_t)(b >> 2) * M + m) * XQ_ROW_BYTES;
    uint4* qdst = reinterpret_cast<uint4*>(rowp + (b & 3) * 32);
    qdst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    qdst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
    reinterpret_cast<float*>(rowp + 128)[b & 3] = __half2float(__float2half_rn(d));
    reinterpret_cast<int*>(rowp + 144)[b & 3] = 8 * sumq;
}

template <int MT>
__global__ void __launch_bounds__(GEMV_THREADS, 1) gemv_q40_kernel(const __grid_constant__ GemvArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    trace_begin(a.tr);
    const int S = a.S, R = a.R, M = a.M;
    const int nk_max = (a.KT + S - 1) / S;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);                               // barriers first (8-byte aligned)
    uint64_t* empty = full + GEMV_NS;
    uint64_t* act_full = empty + GEMV_NS;
    uint64_t* red_bar = act_full + 1;                                                // leader: S-1 remote arrivals
    int* sflag = reinterpret_cast<int*>(red_bar + 1);
    uint8_t* stages = smem + 256;                                                    // [NS][9216]
    uint8_t* act_s = stages + GEMV_NS * Q40T_TILE_BYTES;                             // [nk][M][160]
    float* red = reinterpret_cast<float*>(act_s + (size_t)nk_max * M * XQ_ROW_BYTES);   // [NG*R][MT][rows] (<= NG*MT*128)
    float* out_s = red + GEMV_NG * MT * 128;                                         // [MT][128]
    float* ssq_w = out_s + MT * 128;                                                 // [GEMV_WARPS][MT]
    float* ssq_all = ssq_w + GEMV_WARPS * MT;                                        // [S][MT] filled by all cluster ranks
    float* recv = ssq_all + 8 * MT;                                                  // leader: [S-1][MT][128]

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int crank = (S > 1) ? (int)cg::this_cluster().block_rank() : 0;
    const int cl = blockIdx.x / S;                  // cluster (work unit owner) index
    const int n_owners = gridDim.x / S;
    const int rows = Q40T_TILE_ROWS / R;
    const int kt0 = (a.KT * crank) / S, kt1 = (a.KT * (crank + 1)) / S;
    const int nk = kt1 - kt0;
    // units (row tile, row part) handled by this cluster: cl, cl + n_owners, ...
    const int n_my = (a.n_units - cl + n_owners - 1) / n_owners;
    const int n_it = n_my * nk;
    const bool is_producer = (tid == GEMV_CONSUMERS);
    const uint32_t nib_bytes = rows * 64, sc_bytes = rows * 8;
    const int pre = n_it < GEMV_NS ? n_it : GEMV_NS;
    uint64_t pol = 0;

    // producer lanes: lane p of the producer warp owns iterations it == p (mod GEMV_NPROD); a single
    // issuing thread caps a CTA at ~34 GB/s (measured), four of them lift the cap above the SM's share
    const int rshift = (R == 1) ? 0 : ((R == 2) ? 1 : 2);
    auto issue = [&](int u, int i, int s) {
        const int unit = cl + u * n_owners;
        const int rt = unit >> rshift, rs = unit & (R - 1);
        const uint8_t* t = a.w + ((size_t)rt * a.KT + kt0 + i) * Q40T_TILE_BYTES;
        uint8_t* st = stages + s * Q40T_TILE_BYTES;
        if (R == 1) {
            // nibbles and scales of a whole tile are contiguous: one 9216-byte copy
            mbar_arrive_expect_tx(&full[s], Q40T_TILE_BYTES);
            bulk_g2s_hint(st, t, Q40T_TILE_BYTES, &full[s], pol);
        } else {
            mbar_arrive_expect_tx(&full[s], nib_bytes + sc_bytes);
            bulk_g2s_hint(st, t + rs * nib_bytes, nib_bytes, &full[s], pol);
            bulk_g2s_hint(st + Q40T_NIB_BYTES, t + Q40T_SCALE_OFF + rs * sc_bytes, sc_bytes, &full[s], pol);
        }
    };
    const bool is_prod_lane = (warp == GEMV_CONSUMERS / 32) && (lane < GEMV_NPROD);
    int p_it = lane, p_u = 0, p_i = lane;          // this producer lane's next iteration (it, unit index, k-tile)
    auto p_norm = [&]() { while (p_i >= nk && nk > 0) { p_i -= nk; p_u++; } };

    if (tid == 0) {
        for (int s = 0; s < GEMV_NS; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 4);
        }
        mbar_init(act_full, 1);
        mbar_init(red_bar, S > 1 ? S - 1 : 1);
        mbar_fence_init();
    }
    // cluster-wide: the leader's red_bar must be initialised before any peer can arrive on it
    if (S > 1) cg::this_cluster().sync(); else __syncthreads();
    pdl_launch_dependents();
    GEMV_TRACE(0);

    if (is_prod_lane) {
        // weights do not depend on the previous kernel: stream the first ring-full before the PDL wait
        pol = l2_policy_evict_first();
        p_norm();
        for (; p_it < pre; p_it += GEMV_NPROD, p_i += GEMV_NPROD) {
            p_norm();
            issue(p_u, p_i, p_it);
        }
    }

    if (a.fused) {
        // ---------------- fused prologue: flags wait + merge-add + RMSNorm + Q80 -> act_s ----------------
        const DlNormProlog& nm = a.norm;
        const int dim = a.n;
        const int e0 = kt0 * 128, e1 = kt1 * 128;                 // this CTA's K range (cluster ranks tile the row)
        const int nkb = nk * 4;
        const int items = M * nkb;
        const bool one_pass = items <= GEMV_THREADS;              // a whole 32-block per thread, kept in registers
        // the norm weights are model weights: fetch them before waiting on the previous kernel
        float4 wreg[8];
        if (one_pass && tid < items) {
            const float4* wn = reinterpret_cast<const float4*>(nm.norm_w + e0 + (tid % nkb) * 32);
#pragma unroll
            for (int j = 0; j < 8; j++) wreg[j] = wn[j];
        }
        pdl_wait();
        if (nm.flags != nullptr) {
            if (tid < nm.n_src) {
                const uint32_t epoch = ld_volatile_u32(nm.step_counter) * nm.flag_mul + nm.flag_add;
                wait_flag_ge(nm.flags + tid, epoch);
            }
            __syncthreads();
        }
        const int ll = nm.ll;
        const uint32_t ll_epoch = ll ? ld_volatile_u32(nm.step_counter) * nm.flag_mul + nm.flag_add : 0u;
        const bool writer = (nm.x_out != nullptr) && (cl == 0);
        float ss[MT];
#pragma unroll
        for (int m = 0; m < MT; m++) ss[m] = 0.0f;
        float4 v[8];
        int my_m = 0, my_b = 0;
        if (one_pass) {
            if (tid < items) {
                my_m = tid / nkb;
                my_b = tid % nkb;
                const size_t e = (size_t)my_m * dim + e0 + my_b * 32;
#pragma unroll
                for (int j = 0; j < 8; j++) v[j] = reinterpret_cast<const float4*>(nm.x_in + e)[j];
                for (int s = 0; s < nm.n_src; s++) {
                    const size_t pe = (size_t)s * nm.src_stride + e;
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const float4 t = part_load4(nm.partials, pe + 4 * j, ll, ll_epoch);
                        v[j].x += t.x; v[j].y += t.y; v[j].z += t.z; v[j].w += t.w;
                    }
                }
                float sq = 0.0f;
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (writer) reinterpret_cast<float4*>(nm.x_out + e)[j] = v[j];
                    sq += v[j].x * v[j].x + v[j].y * v[j].y + v[j].z * v[j].z + v[j].w * v[j].w;
                }
#pragma unroll
                for (int m = 0; m < MT; m++) ss[m] = (m == my_m) ? sq : 0.0f;
            }
        } else {
            // pass 1 of 2: coalesced float4 sweep for the sum of squares; row tile 0 publishes x_out
#pragma unroll
            for (int m = 0; m < MT; m++) {
                if (m < M) {
                    const float4* xi = reinterpret_cast<const float4*>(nm.x_in + (size_t)m * dim);
                    for (int i4 = e0 / 4 + tid; i4 < e1 / 4; i4 += GEMV_THREADS) {
                        float4 x4 = xi[i4];
                        for (int s = 0; s < nm.n_src; s++) {
                            const float4 p = part_load4(nm.partials, (size_t)s * nm.src_stride + (size_t)m * dim + 4 * (size_t)i4, ll, ll_epoch);
                            x4.x += p.x; x4.y += p.y; x4.z += p.z; x4.w += p.w;
                        }
                        if (writer) reinterpret_cast<float4*>(nm.x_out + (size_t)m * dim)[i4] = x4;
                        ss[m] += x4.x * x4.x + x4.y * x4.y + x4.z * x4.z + x4.w * x4.w;
                    }
                }
            }
        }
#pragma unroll
        for (int m = 0; m < MT; m++) {
            const float t = warp_sum(ss[m]);
            if (lane == 0) ssq_w[warp * MT + m] = t;
        }
        __syncthreads();
        if (warp < MT) {                                  // warp m reduces row m's per-warp partials
            float t = (lane < GEMV_WARPS) ? ssq_w[lane * MT + warp] : 0.0f;
            t = warp_sum(t);
            const int tidm = warp;
            if (lane == 0) {
            // publish this CTA's share into every cluster rank's table (DSMEM stores): after the
            // barrier nobody reads remote shared memory, so ranks may exit independently later on
            if (S > 1) {
                for (int c = 0; c < S; c++)
                    st_shared_cluster_f32(mapa_shared(smem_u32(ssq_all + crank * MT + tidm), c), t);
            } else {
                ssq_all[tidm] = t;
            }
            }
        }
        if (S > 1) cg::this_cluster().sync(); else __syncthreads();
        GEMV_TRACE(4);
        float inv[MT];
#pragma unroll
        for (int m = 0; m < MT; m++) {
            float t = 0.0f;
            for (int c = 0; c < S; c++) t += ssq_all[c * MT + m];
            inv[m] = 1.0f / sqrtf(t / (float)dim + nm.eps);
        }
        // y = w * (inv * x) -> Q80 -> resident activation slabs, one THREAD per 32-block (no shuffles)
        for (int item = tid; item < items; item += GEMV_THREADS) {
            const int m = one_pass ? my_m : item / nkb, b = one_pass ? my_b : item % nkb;
            if (!one_pass) {
                const size_t e = (size_t)m * dim + e0 + b * 32;
#pragma unroll
                for (int j = 0; j < 8; j++) v[j] = reinterpret_cast<const float4*>(nm.x_in + e)[j];
                for (int s = 0; s < nm.n_src; s++) {
                    const size_t pe = (size_t)s * nm.src_stride + e;
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const float4 t = part_load4(nm.partials, pe + 4 * j, ll, ll_epoch);
                        v[j].x += t.x; v[j].y += t.y; v[j].z += t.z; v[j].w += t.w;
                    }
                }
            }
            float im = inv[0];
#pragma unroll
            for (int q = 1; q < MT; q++) im = (m == q) ? inv[q] : im;
            const float4* wn = reinterpret_cast<const float4*>(nm.norm_w + e0 + b * 32);
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float4 w4 = one_pass ? wreg[j] : wn[j];
                v[j].x = w4.x * (im * v[j].x); v[j].y = w4.y * (im * v[j].y);
                v[j].z = w4.z * (im * v[j].z); v[j].w = w4.w * (im * v[j].w);
            }
            store_q80_block_smem(v, act_s, M, m, b);
        }
        __syncthreads();
        GEMV_TRACE(6);
    }

    if (warp == GEMV_CONSUMERS / 32) {
        // ------------------------------ producer warp ------------------------------
        if (is_producer) {
            if (!a.fused) {
                pdl_wait();
                const uint32_t act_bytes = (uint32_t)nk * M * XQ_ROW_BYTES;
                mbar_arrive_expect_tx(act_full, act_bytes);
                bulk_g2s(act_s, a.xq + (size_t)kt0 * M * XQ_ROW_BYTES, act_bytes, act_full);
            }
        }
        if (is_prod_lane) {
            for (; p_it < n_it; p_it += GEMV_NPROD, p_i += GEMV_NPROD) {
                p_norm();
                const int s = p_it % GEMV_NS;
                mbar_wait(&empty[s], ((p_it / GEMV_NS) & 1) ^ 1);
                issue(p_u, p_i, s);
            }
        }
    } else {
        // ------------------------------ consumers ------------------------------
        const int g = tid >> 7;              // consumer group: k-tile i of a unit goes to group i % NG
        const int t = tid & 127;
        const int row = t % rows;            // local row
        const int kq = t / rows;             // handles blocks kb with (kb % R) == kq
        if (!a.fused) mbar_wait(act_full, 0);
        pdl_wait();                          // already satisfied (producer / prologue waited); orders the epilogue's reads
        if (tid == 0) { GEMV_TRACE(1); }

        for (int u = 0; u < n_my; u++) {
            const int unit = cl + u * n_owners;
            const int rt = unit / R, rs = unit % R;
            const int rtile = rs * rows + row;   // row inside the 128-row tile (for the rotation)
            float acc[MT];
#pragma unroll
            for (int m = 0; m < MT; m++) acc[m] = 0.0f;
            for (int i = g; i < nk; i += GEMV_NG) {
                const int it = u * nk + i;
                const int s = it % GEMV_NS;
                mbar_wait(&full[s], (it / GEMV_NS) & 1);
                const uint8_t* st = stages + s * Q40T_TILE_BYTES;
                const uint8_t* act = act_s + (size_t)i * M * XQ_ROW_BYTES;
                // fully unrolled over the 4 blocks of the tile row (predicated on ownership when R > 1)
                // so the loads and dp4a chains of different blocks overlap
#pragma unroll
                for (int kb = 0; kb < 4; kb++) {
                    if ((kb & (R - 1)) != kq) continue;
                    const uint4 nb = *reinterpret_cast<const uint4*>(st + row * 64 + (((kb + (rtile >> 1)) & 3) << 4));
                    const float dw = __half2float(*reinterpret_cast<const __half*>(st + Q40T_NIB_BYTES + row * 8 + kb * 2));
                    const int lo0 = nb.x & 0x0F0F0F0F, hi0 = (nb.x >> 4) & 0x0F0F0F0F;
                    const int lo1 = nb.y & 0x0F0F0F0F, hi1 = (nb.y >> 4) & 0x0F0F0F0F;
                    const int lo2 = nb.z & 0x0F0F0F0F, hi2 = (nb.z >> 4) & 0x0F0F0F0F;
                    const int lo3 = nb.w & 0x0F0F0F0F, hi3 = (nb.w >> 4) & 0x0F0F0F0F;
#pragma unroll
                    for (int m = 0; m < MT; m++) {
                        if (m < M) {
                            const uint8_t* xr = act + m * XQ_ROW_BYTES;
                            const int4 xl = *reinterpret_cast<const int4*>(xr + kb * 32);
                            const int4 xh = *reinterpret_cast<const int4*>(xr + kb * 32 + 16);
                            // two independent dp4a chains (low / high nibbles) for ILP
                            int is0 = __dp4a(lo0, xl.x, 0), is1 = __dp4a(hi0, xh.x, 0);
                            is0 = __dp4a(lo1, xl.y, is0); is1 = __dp4a(hi1, xh.y, is1);
                            is0 = __dp4a(lo2, xl.z, is0); is1 = __dp4a(hi2, xh.z, is1);
                            is0 = __dp4a(lo3, xl.w, is0); is1 = __dp4a(hi3, xh.w, is1);
                            const float dx = reinterpret_cast<const float*>(xr + 128)[kb];
                            const int s8 = reinterpret_cast<const int*>(xr + 144)[kb];
                            acc[m] += (float)(is0 + is1 - s8) * (dw * dx);
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[s]);
            }
            if (tid == 0 && u == n_my - 1) { GEMV_TRACE(2); }
            // partial sums of the NG*R (group, kq) owners of each row -> out_s
#pragma unroll
            for (int m = 0; m < MT; m++) red[((g * R + kq) * MT + m) * rows + row] = acc[m];
            consumer_bar();
            for (int idx = tid; idx < MT * rows; idx += GEMV_CONSUMERS) {
                const int m = idx / rows, r = idx % rows;
                float vsum = 0.0f;
                for (int j = 0; j < GEMV_NG * R; j++) vsum += red[(j * MT + m) * rows + r];
                out_s[m * 128 + r] = vsum;
            }
            if (S > 1) break;                // split-K: finish below (exactly one unit per cluster)
            consumer_bar();
            epilogue_rows(a.epi, out_s, 128, rows, rt * Q40T_TILE_ROWS + rs * rows, a.d_real, M, 0, tid);
            if (u + 1 < n_my) consumer_bar();   // out_s / red are reused by the next unit
        }
    }
    GEMV_TRACE(3);
    if (S > 1) {
        // split-K reduction without a second cluster barrier: peers push their partial tile into the
        // leader's shared memory (DSMEM stores), release-arrive on its mbarrier and exit; the leader
        // acquires the barrier and sums in rank order (deterministic).
        __syncthreads();
        if (crank != 0) {
            const uint32_t dst = mapa_shared(smem_u32(recv + (size_t)(crank - 1) * MT * 128), 0);
            for (int idx = tid; idx < MT * rows; idx += GEMV_THREADS) {
                const int m = idx / rows, r = idx % rows;
                st_shared_cluster_f32(dst + (uint32_t)(m * 128 + r) * 4u, out_s[m * 128 + r]);
            }
            fence_acq_rel_cluster();
            __syncthreads();
            if (tid == 0) mbar_arrive_remote(mapa_shared(smem_u32(red_bar), 0));
            trace_end(a.tr);
            return;
        }
        mbar_wait_cluster(red_bar, 0);
        for (int idx = tid; idx < MT * rows; idx += GEMV_THREADS) {
            const int m = idx / rows, r = idx % rows;
            float vsum = out_s[m * 128 + r];
            for (int c = 0; c < S - 1; c++) vsum += recv[(size_t)c * MT * 128 + m * 128 + r];
            out_s[m * 128 + r] = vsum;
        }
        __syncthreads();
        pdl_wait();
        const int rt = cl / R, rs = cl % R;
        epilogue_rows(a.epi, out_s, 128, rows, rt * Q40T_TILE_ROWS + rs * rows, a.d_real, M, 0, tid);
    }
    epilogue_complete(a.epi, tid, a.n_epi_ctas, sflag);
    GEMV_TRACE(5);
    trace_end(a.tr);
}

static size_t gemv_smem_bytes(int MT, int nk_max, int M) {
    return 256 + (size_t)GEMV_NS * Q40T_TILE_BYTES + (size_t)nk_max * M * XQ_ROW_BYTES +
           (size_t)(GEMV_NG * MT * 128 + MT * 128) * 4 + (size_t)(GEMV_WARPS * MT + 8 * MT) * 4 +
           (size_t)7 * MT * 128 * 4 + 16;
}

static int mt_of(int M) { return M == 1 ? 1 : (M == 2 ? 2 : (M <= 4 ? 4 : 8)); }

extern "C" long long dl_gemv_q40_smem_bytes(int n, int M, int k_split) {
    const int KT = n / Q40T_TILE_K;
    if (k_split > KT) k_split = 1;
    return (long long)gemv_smem_bytes(mt_of(M), (KT + k_split - 1) / k_split, M);
}

extern void dl_set_error(const char* fmt, ...);
static unsigned long long* g_gemv_trace = nullptr;
// debug: per-CTA timeline (8 x globaltimer ns) written by thread 0 when non-NULL
extern "C" void dl_debug_set_gemv_trace(unsigned long long* p) { g_gemv_trace = p; }

template <int MT>
static int launch_gemv(const GemvArgs& a, int grid, int pdl, cudaStream_t stream) {
    static size_t attr_max = 0;
    const size_t smem = gemv_smem_bytes(MT, (a.KT + a.S - 1) / a.S, a.M);
    if (smem > 227 * 1024) { dl_set_error("dl_gemv_q40: %zu B of shared memory needed (n=%d, M=%d, k_split=%d): raise k_split", smem, a.n, a.M, a.S); return -1; }
    if (smem > attr_max) {
        cudaError_t e = cudaFuncSetAttribute(gemv_q40_kernel<MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { dl_set_error("gemv smem attr: %s", cudaGetErrorString(e)); return -(int)e; }
        attr_max = smem;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
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

// row_split R in {1,2,4}: a 128-row tile is cut into R row parts; k_split S in 1..8: the K range of a
// unit is cut over a cluster of S CTAs; tiles_per_cta T >= 1 (only with S == 1): each CTA streams
// through T units back to back.
extern "C" int dl_gemv_q40(const uint8_t* w_q40t, int d, int n, const uint8_t* xq, int M, int row_split, int k_split,
                           int tiles_per_cta, const DlNormProlog* norm, const DlEpilogue* epi, int pdl, void* stream) {
    if (M < 1 || M > GEMV_MAX_M) { dl_set_error("dl_gemv_q40: M=%d out of range 1..8", M); return -1; }
    if (n % Q40T_TILE_K != 0) { dl_set_error("dl_gemv_q40: n=%d not a multiple of 128", n); return -1; }
    if (row_split != 1 && row_split != 2 && row_split != 4) { dl_set_error("dl_gemv_q40: row_split must be 1,2,4"); return -1; }
    if (k_split < 1 || k_split > 8) { dl_set_error("dl_gemv_q40: k_split must be 1..8"); return -1; }
    if (tiles_per_cta < 1) tiles_per_cta = 1;
    if (epi->kind == DL_EPI_SILU_MUL && (row_split > 2 || d % 64 != 0)) { dl_set_error("dl_gemv_q40: silu epilogue needs row_split<=2, d%%64==0"); return -1; }
    if (norm == nullptr && xq == nullptr) { dl_set_error("dl_gemv_q40: need xq or a norm prologue"); return -1; }
    if (norm != nullptr && norm->n_src > DL_MAX_PEERS) { dl_set_error("dl_gemv_q40: n_src > 8"); return -1; }
    GemvArgs a;
    a.w = w_q40t;
    a.xq = xq;
    a.d_real = d;
    a.d_pad = (d + Q40T_TILE_ROWS - 1) / Q40T_TILE_ROWS * Q40T_TILE_ROWS;
    a.KT = n / Q40T_TILE_K;
    a.n = n;
    if (k_split > a.KT) k_split = 1;
    if (k_split > 1) tiles_per_cta = 1;
    a.M = M;
    a.R = row_split;
    a.S = k_split;
    a.T = tiles_per_cta;
    a.n_units = (a.d_pad / Q40T_TILE_ROWS) * a.R;
    const int owners = (a.n_units + a.T - 1) / a.T;
    a.n_epi_ctas = owners;
    a.fused = norm != nullptr ? 1 : 0;
    if (norm != nullptr) a.norm = *norm; else memset(&a.norm, 0, sizeof(a.norm));
    a.epi = *epi;
    a.trace = g_gemv_trace;
    a.tr = dl_trace_next(TR_GEMV);
    const int grid = owners * a.S;
    cudaStream_t st = (cudaStream_t)stream;
    switch (mt_of(M)) {
        case 1: return launch_gemv<1>(a, grid, pdl, st);
        case 2: return launch_gemv<2>(a, grid, pdl, st);
        case 4: return launch_gemv<4>(a, grid, pdl, st);
        default: return launch_gemv<8>(a, grid, pdl, st);
    }
}
