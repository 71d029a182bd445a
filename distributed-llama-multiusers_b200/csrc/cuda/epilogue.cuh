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
        *reinterpret_cast<__half*>(xh + off) = __float2half_rn((float)q * df);
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
// activation rows m0..m0+M-1; grow0 = global row of local row 0, d_real = real (unpadded) rows.
DL_DEVINL void epilogue_rows(const DlEpilogue& e, const float* out_s, int out_ld, int rows, int grow0, int d_real,
                             int M, int m0, int tid) {
    const int lane = tid & 31;
    if (e.kind == DL_EPI_STORE) {
        if (tid < rows) {
            const int grow = grow0 + tid;
            for (int m = 0; m < M; m++) {
                float v = out_s[m * out_ld + tid];
                if (e.q80_roundtrip) v = q80_roundtrip_warp(v);
                if (grow < d_real) {
                    if (e.residual != nullptr) v += e.residual[(size_t)(m0 + m) * e.residual_ld + grow];
                    const size_t off = (size_t)e.dst_off + (size_t)(m0 + m) * e.ld + grow;
                    for (int p = 0; p < e.n_dst; p++) e.dst[p][off] = v;
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
            const float freq = (region < 2 && grow < d_real) ? e.rope_freqs[hd >> 1] : 0.0f;
            for (int m = 0; m < M; m++) {
                float v = out_s[m * out_ld + tid];
                const int p = e.pos[m0 + m];
                if (region < 2) {
                    float other = __shfl_xor_sync(0xffffffffu, v, 1);
                    float sn, cs;
                    sincosf((float)p * freq, &sn, &cs);
                    // even lane holds x0, odd lane holds x1
                    v = (lane & 1) ? (other * sn + v * cs) : (v * cs - other * sn);
                }
                if (grow >= d_real) {
                    // padded row: nothing to store
                } else if (region == 0) {
                    e.q_out[(size_t)(m0 + m) * e.q_dim + rrow] = v;
                } else {
                    const int page = e.block_table[(size_t)e.slot[m0 + m] * e.max_pages + p / e.page_size];
                    const int kvh = rrow / hs;
                    const size_t off =
                        (((size_t)page * e.n_kv_heads + kvh) * e.page_size + (p % e.page_size)) * hs + hd;
                    void* base = (region == 1) ? e.k_cache : e.v_cache;
                    if (e.kv_f32) reinterpret_cast<float*>(base)[off] = v;
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
            for (int m = 0; m < M; m++) {
                float g = out_s[m * out_ld + 2 * tid];
                float u = out_s[m * out_ld + 2 * tid + 1];
                float h = (e.act_gelu ? act_gelu(g) : act_silu(g)) * u;
                q80_block_store(h, lane, m0 + m, (hrow0 + tid) >> 5, e.out_mcap, e.xq_out, e.xh_out);
            }
        }
    }
}

// epilogue_complete: called by ALL threads of every CTA that ran epilogue_rows (uses __syncthreads).
// DL_EPI_STORE with flags: every CTA fences its (peer) stores, the last one releases the flags.
DL_DEVINL void epilogue_complete(const DlEpilogue& e, int tid, int n_epi_ctas, int* smem_flag) {
    if (e.kind != DL_EPI_STORE || e.flag[0] == nullptr) return;
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
        unsigned prev = atomicAdd(e.cta_counter, 1u);
        *smem_flag = (prev == (unsigned)n_epi_ctas - 1u) ? 1 : 0;
    }
    __syncthreads();
    if (*smem_flag) {
        __threadfence_system();
        if (tid < e.n_dst) {
            uint32_t epoch = ld_volatile_u32(e.step_counter) * e.flag_mul + e.flag_add;
            st_release_sys(e.flag[tid], epoch);
        }
        if (tid == 0) *e.cta_counter = 0u;
    }
}
