// Epilogues shared by the dp4a GEMV and the tcgen05 GEMM.  Both kernels end with an f32 tile
// out_s[m * out_ld + r] (r = local output row, m = activation row) in shared memory; the
// epilogue turns it into what the *next* kernel consumes, so no stand-alone elementwise kernels
// (rope / shift / silu / mul / cast / merge-add) exist on the hot path.
#pragma once
#include "common.cuh"
#include "../../../include/dllama_b200.h"

// ---- Q80 block writers -------------------------------------------------------------------------
// One warp holds one 32-element block (lane = element).  Writes the XQ slab entry and/or the XH
// fp16 entry for activation row m, global K-block index kblock.
DL_DEVINL void q80_block_store(float val, int lane, int m, int kblock, int mcap, uint8_t* xq, uint8_t* xh) {
    float amax = warp_max(fabsf(val));
    float d = amax / 127.0f;
    float id = (d != 0.0f) ? (1.0f / d) : 0.0f;
    int q = __float2int_rn(val * id);
    __half d16 = __float2half_rn(d);
    float df = __half2float(d16);
    int kt = kblock >> 2, kb = kblock & 3;
    if (xq != nullptr) {
        int sumq = warp_sum_i(q);
        uint8_t* row = xq + ((size_t)kt * mcap + m) * XQ_ROW_BYTES;
        reinterpret_cast<int8_t*>(row)[kb * 32 + lane] = (int8_t)q;
        if (lane == 0) {
            reinterpret_cast<float*>(row + 128)[kb] = df;
            reinterpret_cast<int*>(row + 144)[kb] = 8 * sumq;
        }
    }
    if (xh != nullptr) {
        int w, s;
        xh_slot(lane, w, s);
        size_t off = (size_t)kt * (size_t)(mcap >> 3) * XH_ROWGROUP_BYTES + (size_t)(m >> 3) * XH_ROWGROUP_BYTES +
                     (size_t)(kb * 4 + w) * 128 + (size_t)(m & 7) * 16 + (size_t)s * 2;
        // fp16 operand of the tcgen05 path: a block whose scale exceeds 515 (|x| > 65405) saturates instead of becoming inf
        *reinterpret_cast<__half*>(xh + off) = __float2half_rn((float)q * fminf(df, 515.0f));
    }
}

// quantize->dequantize one 32-element block held by a warp
DL_DEVINL float q80_roundtrip_warp(float val) {
    float amax = warp_max(fabsf(val));
    float d = amax / 127.0f;
    float id = (d != 0.0f) ? (1.0f / d) : 0.0f;
    int q = __float2int_rn(val * id);
    return (float)q * __half2float(__float2half_rn(d));
}

DL_DEVINL float act_silu(float x) { return x / (1.0f + expf(-x)); }
DL_DEVINL float act_gelu(float x) {
    return 0.5f * x * (1.0f + tanhf(0.797884560802865f * (x + 0.044715f * x * x * x)));
}

// ---- the epilogue --------------------------------------------------------------------------------
// epilogue_rows: called by (at least) the first `rows` threads (tid = local output row; rows is a
// multiple of 32 so every warp is either fully in or fully out).  out_s[m*out_ld + r] holds
// activation rows m0..m0+M-1 (M <= 32); grow0 = global row of local row 0, d_real = real (unpadded)
// rows.  A caller with several thread groups per tile passes (m_first, m_step) so group k handles
// rows k, k+step, ...; everything a row needs from global memory (position, KV page) is fetched
// once per warp by lane m and broadcast with shuffles, so the per-row loop has no dependent loads.
DL_DEVINL void epilogue_rows(const DlEpilogue& e, const float* out_s, int out_ld, int rows, int grow0, int d_real,
                             int M, int m0, int tid, int m_first = 0, int m_step = 1) {
    const int lane = tid & 31;
    if (e.kind == DL_EPI_STORE) {
        if (tid < rows) {
            const int grow = grow0 + tid;
            const bool live = grow < d_real;
            const bool rt = e.q80_roundtrip != 0;
            const int nd = e.n_dst;
            const size_t ld = (size_t)e.ld;
            const size_t res_ld = (size_t)e.residual_ld;
            const float* res = (e.residual != nullptr && live) ? e.residual + (size_t)m0 * res_ld + grow : nullptr;
            const size_t off0 = (size_t)e.dst_off + (size_t)m0 * ld + grow;
            if (nd == 1) {
                float* d0 = e.dst[0] + off0;
#pragma unroll 4
                for (int m = m_first; m < M; m += m_step) {
                    float v = out_s[m * out_ld + tid];
                    if (rt) v = q80_roundtrip_warp(v);
                    if (live) {
                        if (res != nullptr) v += res[m * res_ld];
                        d0[m * ld] = v;
                    }
                }
            } else if (e.ll == 1) {
                // tensor-parallel push, LL words: value and epoch travel in one 8-byte store per peer; the consumer
                // polls the words themselves, so there is nothing to fence and no flag to release afterwards
                const uint32_t epoch = ld_volatile_u32(e.step_counter) * e.flag_mul + e.flag_add;
                for (int m = m_first; m < M; m += m_step) {
                    float v = out_s[m * out_ld + tid];
                    if (rt) v = q80_roundtrip_warp(v);
                    if (live) {
                        const size_t off = off0 + m * ld;
#pragma unroll 1
                        for (int p = 0; p < nd; p++) ll_store(e.dst[p], off, v, epoch);
                    }
                }
            } else if (e.ll == 2) {
                // tensor-parallel push in the reference's own wire format: Q80 blocks (int8 values + one fp16 scale per
                // 32, src/nn/nn-quants.hpp:64-72) instead of the f32 of their round trip -- 1.06 instead of 4 bytes per
                // value on NVLink, which is what bounds prompt-sized exchanges.  Row R = element offset / ld of the f32
                // layout; a wire row is ld int8 followed by ld/32 fp16 scales.  A warp = one 32-value block.
                const size_t rowb = (size_t)ld + (ld >> 4);
                const size_t row0 = ((size_t)e.dst_off + (size_t)m0 * ld) / ld;
                for (int m = m_first; m < M; m += m_step) {
                    const float v = out_s[m * out_ld + tid];
                    const float amax = warp_max(fabsf(v));
                    const float d = amax / 127.0f;
                    const float id = (d != 0.0f) ? (1.0f / d) : 0.0f;
                    const int q = __float2int_rn(v * id);
                    if (live) {
                        const size_t rb = (row0 + m) * rowb;
#pragma unroll 1
                        for (int p = 0; p < nd; p++) {
                            uint8_t* wrow = reinterpret_cast<uint8_t*>(e.dst[p]) + rb;
                            reinterpret_cast<int8_t*>(wrow)[grow] = (int8_t)q;
                            if (lane == 0) *reinterpret_cast<__half*>(wrow + ld + ((grow >> 5) << 1)) = __float2half_rn(d);
                        }
                    }
                }
            } else {
                for (int m = m_first; m < M; m += m_step) {
                    float v = out_s[m * out_ld + tid];
                    if (rt) v = q80_roundtrip_warp(v);
                    if (live) {
                        if (res != nullptr) v += res[m * res_ld];
                        const size_t off = off0 + m * ld;
#pragma unroll 1
                        for (int p = 0; p < nd; p++) e.dst[p][off] = v;
                    }
                }
            }
        }
    } else if (e.kind == DL_EPI_QKV_ROPE) {
        if (tid < rows) {
            const int grow = grow0 + tid;
            const int hs = e.head_size;
            const int region = (grow < e.q_dim) ? 0 : ((grow < e.q_dim + e.kv_dim) ? 1 : 2);   // warp-uniform
            const int rrow = (region == 0) ? grow : ((region == 1) ? grow - e.q_dim : grow - e.q_dim - e.kv_dim);
            const int hd = rrow % hs;
            const int kvh = rrow / hs;
            const bool live = grow < d_real;
            const float freq = (region < 2 && live) ? e.rope_freqs[hd >> 1] : 0.0f;
            // lane m fetches row m's position and (for k/v rows) its KV page; shuffled to the warp below
            int my_pos = 0, my_page = 0;
            if (lane < M) {
                my_pos = e.pos[m0 + lane];
                if (region != 0) my_page = e.block_table[(size_t)e.slot[m0 + lane] * e.max_pages + my_pos / e.page_size];
            }
            const int page_size = e.page_size, n_kv = e.n_kv_heads, q_dim = e.q_dim;
            float* q_out = e.q_out + (size_t)m0 * q_dim + rrow;
            void* base = (region == 1) ? e.k_cache : e.v_cache;
            const bool kv_f32 = e.kv_f32 != 0;
#pragma unroll 2
            for (int m = m_first; m < M; m += m_step) {
                float v = out_s[m * out_ld + tid];
                const int p = __shfl_sync(0xffffffffu, my_pos, m);
                const int page = __shfl_sync(0xffffffffu, my_page, m);
                if (region < 2) {
                    float other = __shfl_xor_sync(0xffffffffu, v, 1);
                    float sn, cs;
                    sincosf((float)p * freq, &sn, &cs);
                    // even lane holds x0, odd lane holds x1
                    v = (lane & 1) ? (other * sn + v * cs) : (v * cs - other * sn);
                }
                if (!live) {
                    // padded row: nothing to store
                } else if (region == 0) {
                    q_out[(size_t)m * q_dim] = v;
                } else {
                    const size_t off = (((size_t)page * n_kv + kvh) * page_size + (p % page_size)) * hs + hd;
                    if (kv_f32) reinterpret_cast<float*>(base)[off] = v;
                    else reinterpret_cast<__half*>(base)[off] = __float2half_rn(v);
                }
            }
        }
    } else {   // DL_EPI_SILU_MUL: rows interleaved (gate_i, up_i); rows/2 hidden values per tile.
        // thread j < rows/2 owns hidden value j: reads local rows 2j, 2j+1; a warp = one Q80 block.
        const int nh = rows >> 1;
        const int hrow0 = grow0 >> 1;
        // d_real % 64 == 0 is required, so validity is uniform per 32-value block (= per warp)
        if (tid < nh && 2 * (hrow0 + (tid & ~31)) < d_real) {
            const bool gelu = e.act_gelu != 0;
            const int mcap = e.out_mcap;
            uint8_t* xq = e.xq_out;
            uint8_t* xh = e.xh_out;
#pragma unroll 2
            for (int m = m_first; m < M; m += m_step) {
                const float2 gu = *reinterpret_cast<const float2*>(out_s + m * out_ld + 2 * tid);
                float h = (gelu ? act_gelu(gu.x) : act_silu(gu.x)) * gu.y;
                q80_block_store(h, lane, m0 + m, (hrow0 + tid) >> 5, mcap, xq, xh);
            }
        }
    }
}

// epilogue_complete: called by ALL threads of every CTA that ran epilogue_rows (uses __syncthreads).
// DL_EPI_STORE with flags: every CTA fences its (peer) stores, the last one releases the flags.
DL_DEVINL void epilogue_complete(const DlEpilogue& e, int tid, int n_epi_ctas, int* smem_flag) {
    if (e.kind != DL_EPI_STORE || e.flag[0] == nullptr) return;
    // Ordering chain (PTX memory model, cumulativity): every thread's peer stores -> bar.sync ->
    // thread 0: fence.sys + counter atomic -> (observed by the last CTA's thread 0) atomic + fence.sys
    // -> bar.sync -> st.release.sys of the flags.  One thread fences per CTA instead of all of them.
    __syncthreads();
    if (tid == 0) {
        __threadfence_system();
        unsigned prev = atomicAdd(e.cta_counter, 1u);
        const int last = (prev == (unsigned)n_epi_ctas - 1u) ? 1 : 0;
        if (last) {
            __threadfence_system();
            *e.cta_counter = 0u;
        }
        *smem_flag = last;
    }
    __syncthreads();
    if (*smem_flag && tid < e.n_dst) {
        uint32_t epoch = ld_volatile_u32(e.step_counter) * e.flag_mul + e.flag_add;
        st_release_sys(e.flag[tid], epoch);
    }
}
