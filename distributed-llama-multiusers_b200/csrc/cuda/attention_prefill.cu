// Causal attention for PROMPT CHUNKS: segments [begin, end) of batch rows that are consecutive positions
// of one sequence (up to PF_MAX_SEG segments per batch, listed in device memory; all other rows -- decode
// rows, short chunks, padding -- are served by the per-row kernel, which skips the segment rows).  The per-row flash-decode kernel (attention.cu) re-reads the
// whole KV history once per row -- for a 4096-token prompt that is ~1 TB of L2 traffic and 2/3 of the
// time to first token.  Here a CTA owns a tile of 64 query rows x one query head, streams each KV page
// of the sequence ONCE per tile through shared memory, and does both products on the tensor cores
// (mma.sync m16n8k16, fp16 operands, f32 accumulate; flash-attention-2 register layout: the score
// accumulators are re-used as the A fragments of the P*V product).
//
// Reference semantics: OP_MULTIHEAD_ATT (multiheadAtt_F32, src/nn/nn-cpu-ops.cpp:749-784) over the
// positions 0..pos of each row, softmax in f32; output cast to Q80 in the XQ / XH slab layout the wo
// matmul consumes (OP_CAST, src/llm.cpp:296-301).  KV cache: fp16 pages [page][kv_head][32][HS].
// `meta` (device): meta[0] = number of segments, then (begin, end) pairs.
#include <math.h>
#include <mma.h>
#include "epilogue.cuh"

#define PF_QTILE 64
#define PF_WARPS 4
#define PF_THREADS (PF_WARPS * 32)
#define PF_PAGE 32
#define PF_STAGES 2            // KV pages in flight (4 measured slower: 255 -> 283 ms on a 4096-token prompt)
#define PF_MAX_SEG 8

struct PfArgs {
    const float* q;            // [M][n_heads*HS]
    const __half* k_cache;     // [pages][n_kv][32][HS]
    const __half* v_cache;
    const int* pos;
    const int* slot;
    const int* block_table;
    const int* meta;           // meta[0] = n_seg, meta[1+2i], meta[2+2i] = [begin, end) of segment i
    int M, n_heads, n_kv_heads, max_pages;
    uint8_t* xq_out;
    uint8_t* xh_out;
    int out_mcap;
    float* y_out;
    unsigned long long* tr;
};

DL_DEVINL void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
DL_DEVINL void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> DL_DEVINL void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

DL_DEVINL void ldmatrix_x4(uint32_t (&r)[4], const void* smem_row) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(smem_u32(smem_row)));
}
DL_DEVINL void ldmatrix_x4_trans(uint32_t (&r)[4], const void* smem_row) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(smem_u32(smem_row)));
}
// D(16x8, f32) += A(16x16, f16 row) * B(16x8, f16 col)
DL_DEVINL void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
DL_DEVINL uint32_t pack_half2(float lo, float hi) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}

template <int HS>
__global__ void __launch_bounds__(PF_THREADS) attention_prefill_kernel(const __grid_constant__ PfArgs a) {
    constexpr int KS = HS / 16;                 // k-steps of the Q*K^T product
    constexpr int NT_O = HS / 8;                // n-tiles of the output
    constexpr int ROW_H = HS + 8;               // padded smem row (halfs): 16-byte rows land on distinct banks for ldmatrix
    constexpr int PAGE_H = PF_PAGE * ROW_H;
    extern __shared__ __align__(128) uint8_t smem[];
    trace_begin(a.tr);
    __half* ks = reinterpret_cast<__half*>(smem);                         // [STAGES][32][ROW_H]
    __half* vs = ks + PF_STAGES * PAGE_H;                                 // [STAGES][32][ROW_H]
    float* out_s = reinterpret_cast<float*>(smem);                        // [64][HS] f32, reuses the ring after the loop

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, tig = lane & 3;
    const int head = blockIdx.y;
    const int G = a.n_heads / a.n_kv_heads;
    const int kvh = head / G;
    pdl_launch_dependents();
    pdl_wait();                                                            // q, the KV rows of this chunk and meta come from earlier kernels
    // which (segment, 64-row tile) is this CTA?
    int tile = blockIdx.x, begin = 0, end = 0;
    {
        const int n_seg = min(a.meta[0], PF_MAX_SEG);
        int i = 0;
        for (; i < n_seg; i++) {
            begin = a.meta[1 + 2 * i];
            end = a.meta[2 + 2 * i];
            const int nt = (end - begin + PF_QTILE - 1) / PF_QTILE;
            if (tile < nt) break;
            tile -= nt;
        }
        if (i == n_seg) { trace_end(a.tr); return; }
    }
    const int n_rows = end - begin;
    const int r_first = tile * PF_QTILE;                                   // chunk-relative first row of this tile

    const int p0 = a.pos[begin];
    const int slot = a.slot[begin];
    const int* bt = a.block_table + (size_t)slot * a.max_pages;
    const int rows_here = min(PF_QTILE, n_rows - r_first);
    const int last_pos = p0 + r_first + rows_here - 1;
    const int n_pages = last_pos / PF_PAGE + 1;

    // ---- Q fragments (pre-scaled by 1/sqrt(HS)), straight from global f32 ----
    const int rw = r_first + warp * 16;                                    // chunk-relative first row of this warp
    const float scale = rsqrtf((float)HS);
    uint32_t qf[KS][4];
    {
        const int ra = rw + g, rb = rw + g + 8;
        const float* qa = a.q + (size_t)(begin + min(ra, n_rows - 1)) * a.n_heads * HS + head * HS;
        const float* qb = a.q + (size_t)(begin + min(rb, n_rows - 1)) * a.n_heads * HS + head * HS;
#pragma unroll
        for (int k = 0; k < KS; k++) {
            const float2 a0 = *reinterpret_cast<const float2*>(qa + k * 16 + tig * 2);
            const float2 a1 = *reinterpret_cast<const float2*>(qb + k * 16 + tig * 2);
            const float2 a2 = *reinterpret_cast<const float2*>(qa + k * 16 + 8 + tig * 2);
            const float2 a3 = *reinterpret_cast<const float2*>(qb + k * 16 + 8 + tig * 2);
            qf[k][0] = pack_half2(a0.x * scale, a0.y * scale);
            qf[k][1] = pack_half2(a1.x * scale, a1.y * scale);
            qf[k][2] = pack_half2(a2.x * scale, a2.y * scale);
            qf[k][3] = pack_half2(a3.x * scale, a3.y * scale);
        }
    }
    const int pos_a = p0 + rw + g, pos_b = pos_a + 8;                      // positions of this thread's two rows

    float o[NT_O][4];
#pragma unroll
    for (int i = 0; i < NT_O; i++) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.0f;
    float m_a = -INFINITY, m_b = -INFINITY, l_a = 0.0f, l_b = 0.0f;

    auto load_page = [&](int j, int st) {
        const int page = bt[j];
        const __half* kg = a.k_cache + ((size_t)page * a.n_kv_heads + kvh) * PF_PAGE * HS;
        const __half* vg = a.v_cache + ((size_t)page * a.n_kv_heads + kvh) * PF_PAGE * HS;
        constexpr int CH = PF_PAGE * HS / 8;                               // 16-byte chunks per page
        for (int c = tid; c < CH; c += PF_THREADS) {
            const int row = c / (HS / 8), col = (c % (HS / 8)) * 8;
            cp_async16(ks + st * PAGE_H + row * ROW_H + col, kg + row * HS + col);
            cp_async16(vs + st * PAGE_H + row * ROW_H + col, vg + row * HS + col);
        }
        cp_async_commit();
    };

    // PF_STAGES-1 pages ahead; an empty commit keeps one cp.async group per iteration so that
    // wait_group<PF_STAGES-1> always means "page j has landed"
    for (int j = 0; j < PF_STAGES - 1; j++) {
        if (j < n_pages) load_page(j, j); else cp_async_commit();
    }
    for (int j = 0; j < n_pages; j++) {
        const int st = j % PF_STAGES;
        if (j + PF_STAGES - 1 < n_pages) load_page(j + PF_STAGES - 1, (j + PF_STAGES - 1) % PF_STAGES); else cp_async_commit();
        cp_async_wait<PF_STAGES - 1>();
        __syncthreads();
        const __half* kp = ks + st * PAGE_H;
        const __half* vp = vs + st * PAGE_H;
        // tokens past the end of the sequence have never been written: their scores are masked to -inf, but a
        // zero probability times a non-finite V would still poison the product -> clear those V rows
        const int valid = p0 + n_rows - j * PF_PAGE;                       // tokens of this page that exist
        if (valid < PF_PAGE) {
            __half* vz = vs + st * PAGE_H;
            for (int c = tid; c < (PF_PAGE - valid) * (HS / 8); c += PF_THREADS)
                *reinterpret_cast<uint4*>(vz + (valid + c / (HS / 8)) * ROW_H + (c % (HS / 8)) * 8) = make_uint4(0, 0, 0, 0);
            __syncthreads();
        }
        // warps whose rows all lie before this page have nothing to add (causal)
        if (j * PF_PAGE <= p0 + rw + 15) {
            // ---- S = Q K^T : 16 rows x 32 tokens ----
            float s[4][4];
#pragma unroll
            for (int nt = 0; nt < 4; nt++) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.0f;
#pragma unroll
            for (int nt = 0; nt < 4; nt++) {
#pragma unroll
                for (int k2 = 0; k2 < KS / 2; k2++) {
                    uint32_t b[4];
                    ldmatrix_x4(b, kp + (nt * 8 + (lane & 7)) * ROW_H + k2 * 32 + (lane >> 3) * 8);
                    mma16816(s[nt], qf[2 * k2], b[0], b[1]);
                    mma16816(s[nt], qf[2 * k2 + 1], b[2], b[3]);
                }
            }
            // ---- causal mask + online softmax (rows g and g+8 of the warp tile) ----
            float mx_a = -INFINITY, mx_b = -INFINITY;
#pragma unroll
            for (int nt = 0; nt < 4; nt++) {
                const int t0 = j * PF_PAGE + nt * 8 + tig * 2;
                if (t0 > pos_a) s[nt][0] = -INFINITY;
                if (t0 + 1 > pos_a) s[nt][1] = -INFINITY;
                if (t0 > pos_b) s[nt][2] = -INFINITY;
                if (t0 + 1 > pos_b) s[nt][3] = -INFINITY;
                mx_a = fmaxf(mx_a, fmaxf(s[nt][0], s[nt][1]));
                mx_b = fmaxf(mx_b, fmaxf(s[nt][2], s[nt][3]));
            }
            mx_a = fmaxf(mx_a, __shfl_xor_sync(0xffffffffu, mx_a, 1));
            mx_a = fmaxf(mx_a, __shfl_xor_sync(0xffffffffu, mx_a, 2));
            mx_b = fmaxf(mx_b, __shfl_xor_sync(0xffffffffu, mx_b, 1));
            mx_b = fmaxf(mx_b, __shfl_xor_sync(0xffffffffu, mx_b, 2));
            const float mn_a = fmaxf(m_a, mx_a), mn_b = fmaxf(m_b, mx_b);
            // a row with nothing visible yet keeps m = -inf: use 0 as the reference so exp() stays finite
            const float ref_a = (mn_a == -INFINITY) ? 0.0f : mn_a, ref_b = (mn_b == -INFINITY) ? 0.0f : mn_b;
            const float sc_a = __expf(m_a - ref_a), sc_b = __expf(m_b - ref_b);      // exp(-inf) = 0 on the first page
            m_a = mn_a; m_b = mn_b;
            l_a *= sc_a; l_b *= sc_b;
            uint32_t pf[2][4];
#pragma unroll
            for (int nt = 0; nt < 4; nt++) {
                const float e0 = __expf(s[nt][0] - ref_a), e1 = __expf(s[nt][1] - ref_a);
                const float e2 = __expf(s[nt][2] - ref_b), e3 = __expf(s[nt][3] - ref_b);
                l_a += e0 + e1;
                l_b += e2 + e3;
                pf[nt >> 1][(nt & 1) * 2 + 0] = pack_half2(e0, e1);
                pf[nt >> 1][(nt & 1) * 2 + 1] = pack_half2(e2, e3);
            }
#pragma unroll
            for (int i = 0; i < NT_O; i++) { o[i][0] *= sc_a; o[i][1] *= sc_a; o[i][2] *= sc_b; o[i][3] *= sc_b; }
            // ---- O += P V : k = 32 tokens (2 k-steps), n = HS ----
#pragma unroll
            for (int kk = 0; kk < 2; kk++) {
#pragma unroll
                for (int on = 0; on < NT_O; on += 2) {
                    uint32_t b[4];
                    ldmatrix_x4_trans(b, vp + (kk * 16 + ((lane >> 3) & 1) * 8 + (lane & 7)) * ROW_H + (on + (lane >> 4)) * 8);
                    mma16816(o[on], pf[kk], b[0], b[1]);
                    mma16816(o[on + 1], pf[kk], b[2], b[3]);
                }
            }
        }
        __syncthreads();                       // everyone is done with stage st before it is refilled
    }

    // ---- finish: divide by the row sums, stage the tile in smem, Q80-cast into the slabs ----
    l_a += __shfl_xor_sync(0xffffffffu, l_a, 1);
    l_a += __shfl_xor_sync(0xffffffffu, l_a, 2);
    l_b += __shfl_xor_sync(0xffffffffu, l_b, 1);
    l_b += __shfl_xor_sync(0xffffffffu, l_b, 2);
    const float inv_a = l_a > 0.0f ? 1.0f / l_a : 0.0f, inv_b = l_b > 0.0f ? 1.0f / l_b : 0.0f;
    {
        float* oa = out_s + (warp * 16 + g) * HS;
        float* ob = out_s + (warp * 16 + g + 8) * HS;
#pragma unroll
        for (int i = 0; i < NT_O; i++) {
            *reinterpret_cast<float2*>(oa + i * 8 + tig * 2) = make_float2(o[i][0] * inv_a, o[i][1] * inv_a);
            *reinterpret_cast<float2*>(ob + i * 8 + tig * 2) = make_float2(o[i][2] * inv_b, o[i][3] * inv_b);
        }
    }
    __syncwarp();                              // a warp only re-reads the 16 rows it wrote
    for (int rr = 0; rr < 16; rr++) {
        const int rc = rw + rr;                // chunk-relative row
        if (rc >= n_rows) break;
        const int m = begin + rc;
#pragma unroll
        for (int b = 0; b < HS / 32; b++) {
            const float v = out_s[(warp * 16 + rr) * HS + b * 32 + lane];
            q80_block_store(v, lane, m, (head * HS) / 32 + b, a.out_mcap, a.xq_out, a.xh_out);
            if (a.y_out != nullptr) a.y_out[(size_t)m * a.n_heads * HS + head * HS + b * 32 + lane] = v;
        }
    }
    trace_end(a.tr);
}

extern void dl_set_error(const char* fmt, ...);

template <int HS>
static int launch_pf(const PfArgs& a, int pdl, cudaStream_t stream) {
    const size_t ring = (size_t)PF_STAGES * 2 * PF_PAGE * (HS + 8) * sizeof(__half);
    const size_t outb = (size_t)PF_QTILE * HS * sizeof(float);
    const size_t smem = ring > outb ? ring : outb;
    static bool set = false;
    if (!set) {
        cudaError_t e = cudaFuncSetAttribute(attention_prefill_kernel<HS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { dl_set_error("attention_prefill smem attr: %s", cudaGetErrorString(e)); return -(int)e; }
        set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((a.M + PF_QTILE - 1) / PF_QTILE + PF_MAX_SEG), (unsigned)a.n_heads);      // upper bound of the tiles of all segments
    cfg.blockDim = dim3(PF_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    if (pdl) {
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
    }
    cudaError_t e = cudaLaunchKernelEx(&cfg, attention_prefill_kernel<HS>, a);
    if (e != cudaSuccess) { dl_set_error("attention_prefill launch: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}

extern "C" int dl_attention_prefill(const float* q, int M, int n_heads, int n_kv_heads, int head_size, const void* k_cache,
                                    const void* v_cache, const int* pos, const int* slot, const int* block_table,
                                    int max_pages, int page_size, const int* meta, uint8_t* xq_out, uint8_t* xh_out,
                                    int out_mcap, float* y_out, int pdl, void* stream) {
    if (page_size != PF_PAGE) { dl_set_error("dl_attention_prefill: page_size must be %d", PF_PAGE); return -1; }
    if (n_heads % n_kv_heads != 0) { dl_set_error("dl_attention_prefill: bad head grouping"); return -1; }
    PfArgs a;
    a.q = q; a.k_cache = (const __half*)k_cache; a.v_cache = (const __half*)v_cache; a.pos = pos; a.slot = slot;
    a.block_table = block_table; a.meta = meta; a.M = M; a.n_heads = n_heads; a.n_kv_heads = n_kv_heads;
    a.max_pages = max_pages; a.xq_out = xq_out; a.xh_out = xh_out; a.out_mcap = out_mcap; a.y_out = y_out;
    a.tr = dl_trace_next(TR_ATT);
    if (head_size == 128) return launch_pf<128>(a, pdl, (cudaStream_t)stream);
    if (head_size == 64) return launch_pf<64>(a, pdl, (cudaStream_t)stream);
    dl_set_error("dl_attention_prefill: unsupported head_size %d (64 or 128)", head_size);
    return -1;
}
