// Paged GQA attention for decode rows and prefill-chunk rows (flash-decode, split-KV).
//
// Replaces OP_MULTIHEAD_ATT (multiheadAtt_F32, src/nn/nn-cpu-ops.cpp:749-784), which materialises
// a [seqLen] score row per head over ONE shared contiguous KV cache.  Here:
//  * every batch row m has its own position pos[m] and its own KV pages (block_table[slot[m]]),
//    so concurrent users never touch each other's cache (fixes reference defects D2/D3);
//  * grid = (split, kv_head, row): each CTA streams a page range of ONE kv head through a
//    4-stage bulk-async-copy ring (a page of one head is contiguous: page_size*head_size
//    elements) and serves all G = n_heads/n_kv_heads query heads of the group from it;
//  * online softmax; scores of 8 tokens x 4 heads are reduced across the warp with a 31-shuffle
//    transpose-reduce, leaving exactly one (token, head) score per lane -> one exp per lane;
//  * split results go to a scratch buffer; the last CTA to finish a (row, kv head) merges them
//    (self-resetting counter), so the whole op is one launch;
//  * the output is cast to Q80 and written directly in the XQ / XH slab layout the wo matmul
//    consumes (OP_CAST y->yq_slice, src/llm.cpp:296-301).
#include <math.h>
#include "epilogue.cuh"

#define ATT_PAGE 32            // tokens per KV page (fixed)
#define ATT_STAGES 4
#define ATT_MIN_SPLIT_PAGES 24    // contexts up to 768 tokens: one CTA per (row, kv head), no split
#define ATT_WARPS 4
#define ATT_THREADS (ATT_WARPS * 32 + 32)
#define ATT_COUNTER_BYTES 65536   // completion counters live at scratch offset 0; caller zeroes scratch once

struct AttArgs {
    const float* q;            // [M][n_heads*HS]
    const void* k_cache;       // [pages][n_kv][ATT_PAGE][HS]
    const void* v_cache;
    const int* pos;
    const int* slot;
    const int* block_table;
    const int* row_skip;       // NULL, or row_skip[m] != 0: row m is served by the prefill kernel
    int M, n_heads, n_kv_heads, max_pages, n_splits;
    unsigned* counters;        // scratch: [M][n_kv_heads] at offset 0 (fixed 64 KB region, stays zero)
    float* part_o;             //          [M][n_heads][n_splits][HS]
    float* part_ml;            //          [M][n_heads][n_splits][2]
    uint8_t* xq_out;
    uint8_t* xh_out;
    int out_mcap;
    float* y_out;              // [M][n_heads*HS] or NULL
    unsigned long long* tr;
};

// 32 values per lane -> lane i ends with the warp-wide sum of value i (31 shuffles)
DL_DEVINL float warp_transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const bool upper = (lane & o) != 0;
#pragma unroll
        for (int k = 0; k < o; k++) {
            const float send = upper ? v[k] : v[k + o];
            const float keep = upper ? v[k + o] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return v[0];
}

template <typename T> DL_DEVINL float kv_to_float(T v);
template <> DL_DEVINL float kv_to_float<float>(float v) { return v; }
template <> DL_DEVINL float kv_to_float<__half>(__half v) { return __half2float(v); }

template <int HS, int NCH, typename KV>
__global__ void __launch_bounds__(ATT_THREADS, 1) attention_kernel(const __grid_constant__ AttArgs a) {
    constexpr int DPL = HS / 32;                       // head dims per lane
    constexpr int PAGE_BYTES = ATT_PAGE * HS * (int)sizeof(KV);
    constexpr int GMAX = NCH * 4;
    extern __shared__ __align__(128) uint8_t smem[];
    trace_begin(a.tr);
    uint8_t* ring = smem;                                                   // [STAGES][2][PAGE_BYTES]
    float* q_s = reinterpret_cast<float*>(ring + ATT_STAGES * 2 * PAGE_BYTES);    // [GMAX][HS] pre-scaled
    float* w_acc = q_s + GMAX * HS;                                         // [WARPS][GMAX][HS]
    float* w_ml = w_acc + ATT_WARPS * GMAX * HS;                            // [WARPS][GMAX][2]
    float* c_ml = w_ml + ATT_WARPS * GMAX * 2;                              // [GMAX][2] CTA-level (M, L)
    uint64_t* full = reinterpret_cast<uint64_t*>(c_ml + GMAX * 2);
    uint64_t* empty = full + ATT_STAGES;
    int* s_last = reinterpret_cast<int*>(empty + ATT_STAGES);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int sp = blockIdx.x, kvh = blockIdx.y, m = blockIdx.z;
    const int G = a.n_heads / a.n_kv_heads;

    if (tid == 0) {
        for (int s = 0; s < ATT_STAGES; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], ATT_WARPS);
        }
        mbar_fence_init();
    }
    __syncthreads();
    pdl_launch_dependents();
    pdl_wait();     // q, pos, block tables and the KV cache all come from earlier kernels
    if (a.row_skip != nullptr && a.row_skip[m] != 0) { trace_end(a.tr); return; }      // uniform for the CTA

    const int T = a.pos[m] + 1;                                // attend to positions 0..pos
    const int n_pages = (T + ATT_PAGE - 1) / ATT_PAGE;
    // The grid carries a.n_splits CTAs per (row, kv head) -- sized for the longest context the engine allows --
    // but the split + merge round trip (scratch, counter, last-CTA merge) costs about as much as streaming 12
    // pages, so a row only uses as many splits as its OWN context justifies; the other CTAs leave at once.
    // (measured, 8B decode: ~170-token context 574 tok/s with 4 splits, 603 with 1; 4096-token context 434 tok/s
    // with 18 splits of 7 pages, 454 with 16 of 8)
    const int n_eff = (n_pages <= ATT_MIN_SPLIT_PAGES) ? 1 : min(a.n_splits, (n_pages + 7) / 8);
    if (sp >= n_eff) { trace_end(a.tr); return; }
    const int pps = (n_pages + n_eff - 1) / n_eff;             // pages per split
    const int pg0 = sp * pps;
    const int pg1 = min(n_pages, pg0 + pps);
    const int np = max(0, pg1 - pg0);
    const int* bt = a.block_table + (size_t)a.slot[m] * a.max_pages;

    if (warp == ATT_WARPS) {
        // producer warp: the 32 lanes fetch 32 block-table entries at once (one load latency instead of one per
        // page in front of every copy), lane 0 issues the bulk copies
        for (int i0 = 0; i0 < np; i0 += 32) {
            const int my_page = (i0 + lane < np) ? bt[pg0 + i0 + lane] : 0;
            const int cnt = min(32, np - i0);
            for (int j = 0; j < cnt; j++) {
                const int page = __shfl_sync(0xffffffffu, my_page, j);
                if (lane == 0) {
                    const int i = i0 + j;
                    const int s = i % ATT_STAGES;
                    if (i >= ATT_STAGES) mbar_wait(&empty[s], ((i / ATT_STAGES) & 1) ^ 1);
                    const size_t poff = ((size_t)page * a.n_kv_heads + kvh) * (size_t)PAGE_BYTES;
                    mbar_arrive_expect_tx(&full[s], 2 * PAGE_BYTES);
                    bulk_g2s(ring + (s * 2 + 0) * PAGE_BYTES, reinterpret_cast<const uint8_t*>(a.k_cache) + poff, PAGE_BYTES, &full[s]);
                    bulk_g2s(ring + (s * 2 + 1) * PAGE_BYTES, reinterpret_cast<const uint8_t*>(a.v_cache) + poff, PAGE_BYTES, &full[s]);
                }
                __syncwarp();
            }
        }
    } else {
        // stage q (pre-scaled by 1/sqrt(HS)); heads >= G are zero
        const float scale = rsqrtf((float)HS);
        for (int i = tid; i < GMAX * HS; i += ATT_WARPS * 32) {
            const int h = i / HS, d = i % HS;
            q_s[i] = (h < G) ? a.q[(size_t)m * a.n_heads * HS + (size_t)(kvh * G + h) * HS + d] * scale : 0.0f;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(ATT_WARPS * 32));

        // running state: lane i owns (m, l) of head (i & 3) of each chunk (replicated over 8 lanes)
        float run_m[NCH], run_l[NCH], acc[NCH][4][DPL];
#pragma unroll
        for (int c = 0; c < NCH; c++) {
            run_m[c] = -INFINITY;
            run_l[c] = 0.0f;
#pragma unroll
            for (int h = 0; h < 4; h++)
#pragma unroll
                for (int d = 0; d < DPL; d++) acc[c][h][d] = 0.0f;
        }
        const int hsel = lane & 3, jsel = lane >> 2;

        for (int i = 0; i < np; i++) {
            const int s = i % ATT_STAGES;
            mbar_wait(&full[s], (i / ATT_STAGES) & 1);
            const KV* Kp = reinterpret_cast<const KV*>(ring + (s * 2 + 0) * PAGE_BYTES);
            const KV* Vp = reinterpret_cast<const KV*>(ring + (s * 2 + 1) * PAGE_BYTES);
            const int ntok = min(ATT_PAGE, T - (pg0 + i) * ATT_PAGE);   // valid tokens in this page
#pragma unroll
            for (int c = 0; c < NCH; c++) {
                float qv[4][DPL];
#pragma unroll
                for (int h = 0; h < 4; h++)
#pragma unroll
                    for (int d = 0; d < DPL; d++) qv[h][d] = q_s[(c * 4 + h) * HS + lane * DPL + d];
                float sc[32];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const int t = warp * 8 + j;
                    float kf[DPL];
#pragma unroll
                    for (int d = 0; d < DPL; d++) kf[d] = kv_to_float<KV>(Kp[t * HS + lane * DPL + d]);
#pragma unroll
                    for (int h = 0; h < 4; h++) {
                        float x = 0.0f;
#pragma unroll
                        for (int d = 0; d < DPL; d++) x += qv[h][d] * kf[d];
                        sc[j * 4 + h] = x;
                    }
                }
                float sv = warp_transpose_reduce32(sc, lane);      // score of (token jsel, head hsel)
                const bool valid = (warp * 8 + jsel < ntok) && (c * 4 + hsel < G);
                sv = valid ? sv : -INFINITY;
                float mx = sv;
                mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
                mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
                mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
                const float m_new = fmaxf(run_m[c], mx);
                const float m_use = (m_new == -INFINITY) ? 0.0f : m_new;
                const float corr = __expf(run_m[c] - m_use);       // run_m = -inf -> 0
                const float p = valid ? __expf(sv - m_use) : 0.0f;
                float ps = p;
                ps += __shfl_xor_sync(0xffffffffu, ps, 4);
                ps += __shfl_xor_sync(0xffffffffu, ps, 8);
                ps += __shfl_xor_sync(0xffffffffu, ps, 16);
                run_l[c] = run_l[c] * corr + ps;
                run_m[c] = m_new;
#pragma unroll
                for (int h = 0; h < 4; h++) {
                    const float ch = __shfl_sync(0xffffffffu, corr, h);
#pragma unroll
                    for (int d = 0; d < DPL; d++) acc[c][h][d] *= ch;
                }
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const int t = warp * 8 + j;
                    float vf[DPL];
#pragma unroll
                    for (int d = 0; d < DPL; d++) vf[d] = (t < ntok) ? kv_to_float<KV>(Vp[t * HS + lane * DPL + d]) : 0.0f;
#pragma unroll
                    for (int h = 0; h < 4; h++) {
                        const float ph = __shfl_sync(0xffffffffu, p, j * 4 + h);
#pragma unroll
                        for (int d = 0; d < DPL; d++) acc[c][h][d] += ph * vf[d];
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
        // per-warp partials -> smem
#pragma unroll
        for (int c = 0; c < NCH; c++) {
#pragma unroll
            for (int h = 0; h < 4; h++)
#pragma unroll
                for (int d = 0; d < DPL; d++) w_acc[(warp * GMAX + c * 4 + h) * HS + lane * DPL + d] = acc[c][h][d];
            if (lane < 4) {
                w_ml[(warp * GMAX + c * 4 + lane) * 2 + 0] = run_m[c];
                w_ml[(warp * GMAX + c * 4 + lane) * 2 + 1] = run_l[c];
            }
        }
    }
    __syncthreads();

    // CTA-level merge of the 4 warps: thread d < HS owns head dim d
    const bool single = (n_eff == 1);
    if (tid < HS) {
        for (int h = 0; h < G; h++) {
            float Mx = -INFINITY;
            for (int w = 0; w < ATT_WARPS; w++) Mx = fmaxf(Mx, w_ml[(w * GMAX + h) * 2]);
            const float Mu = (Mx == -INFINITY) ? 0.0f : Mx;
            float L = 0.0f, O = 0.0f;
            for (int w = 0; w < ATT_WARPS; w++) {
                const float e = __expf(w_ml[(w * GMAX + h) * 2] - Mu);
                L += w_ml[(w * GMAX + h) * 2 + 1] * e;
                O += w_acc[(w * GMAX + h) * HS + tid] * e;
            }
            const int head = kvh * G + h;
            if (single) {
                const float y = (L > 0.0f) ? O / L : 0.0f;
                if (a.y_out != nullptr) a.y_out[(size_t)m * a.n_heads * HS + (size_t)head * HS + tid] = y;
                if (a.xq_out != nullptr || a.xh_out != nullptr)
                    q80_block_store(y, lane, m, (head * HS + tid) >> 5, a.out_mcap, a.xq_out, a.xh_out);
            } else {
                const size_t pidx = ((size_t)m * a.n_heads + head) * a.n_splits + sp;
                a.part_o[pidx * HS + tid] = O;
                if (tid == 0) {
                    a.part_ml[pidx * 2 + 0] = Mx;
                    a.part_ml[pidx * 2 + 1] = L;
                }
            }
        }
    }
    if (single) { trace_end(a.tr); return; }

    // last CTA of this (row, kv head) merges the splits
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned prev = atomicAdd(&a.counters[m * a.n_kv_heads + kvh], 1u);
        *s_last = (prev == (unsigned)n_eff - 1u) ? 1 : 0;
        if (*s_last) a.counters[m * a.n_kv_heads + kvh] = 0u;
    }
    __syncthreads();
    if (!*s_last) { trace_end(a.tr); return; }
    __threadfence();
    // merge: warp w takes heads w, w+4, ...; lane s holds split s's (max, sum) so the rescale factors come out of
    // two warp reductions, then every lane sums its head dims over the splits (independent loads, one latency)
    if (warp < ATT_WARPS) {
        for (int h = warp; h < G; h += ATT_WARPS) {
            const int head = kvh * G + h;
            const size_t base = ((size_t)m * a.n_heads + head) * a.n_splits;
            float ms = -INFINITY, ls = 0.0f;
            if (lane < n_eff) {
                ms = __ldcg(&a.part_ml[(base + lane) * 2]);
                ls = __ldcg(&a.part_ml[(base + lane) * 2 + 1]);
            }
            const float Mx = warp_max(ms);
            const float Mu = (Mx == -INFINITY) ? 0.0f : Mx;
            const float e = (lane < n_eff) ? __expf(ms - Mu) : 0.0f;
            const float L = warp_sum(ls * e);
            float O[DPL];
#pragma unroll
            for (int d = 0; d < DPL; d++) O[d] = 0.0f;
            for (int s2 = 0; s2 < n_eff; s2++) {
                const float es = __shfl_sync(0xffffffffu, e, s2);
#pragma unroll
                for (int d = 0; d < DPL; d++) O[d] += __ldcg(&a.part_o[(base + s2) * HS + lane * DPL + d]) * es;
            }
            const float inv = (L > 0.0f) ? 1.0f / L : 0.0f;
#pragma unroll
            for (int d = 0; d < DPL; d++) w_acc[h * HS + lane * DPL + d] = O[d] * inv;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(ATT_WARPS * 32));
        if (tid < HS) {
            for (int h = 0; h < G; h++) {
                const int head = kvh * G + h;
                const float y = w_acc[h * HS + tid];
                if (a.y_out != nullptr) a.y_out[(size_t)m * a.n_heads * HS + (size_t)head * HS + tid] = y;
                if (a.xq_out != nullptr || a.xh_out != nullptr)
                    q80_block_store(y, lane, m, (head * HS + tid) >> 5, a.out_mcap, a.xq_out, a.xh_out);
            }
        }
    }
    trace_end(a.tr);
}

extern void dl_set_error(const char* fmt, ...);

static size_t att_smem_bytes(int HS, int NCH, int esize) {
    const int GMAX = NCH * 4;
    return (size_t)ATT_STAGES * 2 * ATT_PAGE * HS * esize + (size_t)GMAX * HS * 4 + (size_t)ATT_WARPS * GMAX * HS * 4 +
           (size_t)ATT_WARPS * GMAX * 2 * 4 + (size_t)GMAX * 2 * 4 + 2 * ATT_STAGES * 8 + 16;
}

template <int HS, int NCH, typename KV>
static int launch_att(const AttArgs& a, int pdl, cudaStream_t stream) {
    static bool attr_set = false;
    const size_t smem = att_smem_bytes(HS, NCH, (int)sizeof(KV));
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(attention_kernel<HS, NCH, KV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { dl_set_error("attention smem attr: %s", cudaGetErrorString(e)); return -(int)e; }
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(a.n_splits, a.n_kv_heads, a.M);
    cfg.blockDim = dim3(ATT_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    if (pdl) {
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
    }
    cudaError_t e = cudaLaunchKernelEx(&cfg, attention_kernel<HS, NCH, KV>, a);
    if (e != cudaSuccess) { dl_set_error("attention launch: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}

template <int HS, typename KV>
static int launch_att_nch(const AttArgs& a, int G, int pdl, cudaStream_t stream) {
    if (G <= 4) return launch_att<HS, 1, KV>(a, pdl, stream);
    if (G <= 8) return launch_att<HS, 2, KV>(a, pdl, stream);
    return launch_att<HS, 4, KV>(a, pdl, stream);
}

extern "C" long long dl_attention_scratch_bytes(int M, int n_heads, int head_size, int max_splits) {
    long long o = (long long)M * n_heads * max_splits * head_size * 4;
    long long ml = (long long)M * n_heads * max_splits * 2 * 4;
    return ATT_COUNTER_BYTES + ((o + 255) / 256 + (ml + 255) / 256) * 256;
}

extern "C" int dl_attention(const float* q, int M, int n_heads, int n_kv_heads, int head_size, const void* k_cache,
                            const void* v_cache, int kv_f32, const int* pos, const int* slot, const int* block_table,
                            int max_pages, int page_size, int n_splits, void* scratch, uint8_t* xq_out,
                            uint8_t* xh_out, int out_mcap, float* y_out, const int* row_skip, int pdl, void* stream) {
    if (page_size != ATT_PAGE) { dl_set_error("dl_attention: page_size must be %d", ATT_PAGE); return -1; }
    if (n_heads % n_kv_heads != 0 || n_heads / n_kv_heads > 16) { dl_set_error("dl_attention: unsupported head grouping %d/%d", n_heads, n_kv_heads); return -1; }
    if (n_splits < 1) n_splits = 1;
    if (n_splits > 1 && scratch == nullptr) { dl_set_error("dl_attention: n_splits>1 needs scratch"); return -1; }
    AttArgs a;
    a.q = q; a.k_cache = k_cache; a.v_cache = v_cache; a.pos = pos; a.slot = slot; a.block_table = block_table;
    a.row_skip = row_skip;
    a.M = M; a.n_heads = n_heads; a.n_kv_heads = n_kv_heads; a.max_pages = max_pages; a.n_splits = n_splits;
    // the completion counters are only touched by rows that are split
    if (n_splits > 1 && (long long)M * n_kv_heads * 4 > ATT_COUNTER_BYTES) { dl_set_error("dl_attention: M*n_kv_heads too large"); return -1; }
    long long o = ((long long)M * n_heads * n_splits * head_size * 4 + 255) / 256 * 256;
    a.counters = reinterpret_cast<unsigned*>(scratch);
    a.part_o = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(scratch) + ATT_COUNTER_BYTES);
    a.part_ml = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(scratch) + ATT_COUNTER_BYTES + o);
    a.xq_out = xq_out; a.xh_out = xh_out; a.out_mcap = out_mcap; a.y_out = y_out;
    a.tr = dl_trace_next(TR_ATT);
    const int G = n_heads / n_kv_heads;
    cudaStream_t st = (cudaStream_t)stream;
    if (kv_f32) {
        if (head_size == 128) return launch_att_nch<128, float>(a, G, pdl, st);
        if (head_size == 64) return launch_att_nch<64, float>(a, G, pdl, st);
        if (head_size == 32) return launch_att_nch<32, float>(a, G, pdl, st);
    } else {
        if (head_size == 128) return launch_att_nch<128, __half>(a, G, pdl, st);
        if (head_size == 64) return launch_att_nch<64, __half>(a, G, pdl, st);
        if (head_size == 32) return launch_att_nch<32, __half>(a, G, pdl, st);
    }
    dl_set_error("dl_attention: unsupported head_size %d", head_size);
    return -1;
}
