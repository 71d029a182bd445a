// Weight preparation, activation producers (reduce + RMSNorm + Q80), embedding, argmax,
// error plumbing.
#include <stdarg.h>
#include <stdio.h>
#include "epilogue.cuh"

static thread_local char g_err[512] = "";
void dl_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
extern "C" const char* dl_last_error(void) { return g_err; }

// ---- optional launch timeline -----------------------------------------------------------------
static unsigned long long* g_trace_base = nullptr;
static int g_trace_cap = 0, g_trace_n = 0;
static int g_trace_kind[8192];
unsigned long long* dl_trace_next(int kind) {
    if (g_trace_base == nullptr || g_trace_n >= g_trace_cap || g_trace_n >= 8192) return nullptr;
    g_trace_kind[g_trace_n] = kind;
    return g_trace_base + 2 * (size_t)(g_trace_n++);
}
// buf: device array of 2*cap uint64, pre-filled by the caller with {UINT64_MAX, 0} pairs; NULL disables
extern "C" void dl_trace_set(unsigned long long* buf, int cap) { g_trace_base = buf; g_trace_cap = cap; g_trace_n = 0; }
extern "C" int dl_trace_count(void) { return g_trace_n; }
extern "C" int dl_trace_kind(int i) { return (i >= 0 && i < g_trace_n) ? g_trace_kind[i] : -1; }

#define CHECK_LAUNCH(name)                                                              \
    do {                                                                                \
        cudaError_t e__ = cudaGetLastError();                                           \
        if (e__ != cudaSuccess) {                                                       \
            dl_set_error("%s: %s", name, cudaGetErrorString(e__));                      \
            return -(int)e__;                                                           \
        }                                                                               \
    } while (0)

extern "C" int dl_device_sm_count(void) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
    return n;
}

// ------------------------------------------------------------------------------------------------
// Q40 file rows -> Q40T tiles
// ------------------------------------------------------------------------------------------------
extern "C" long long dl_q40t_bytes(int d, int n) {
    long long d_pad = ((long long)d + Q40T_TILE_ROWS - 1) / Q40T_TILE_ROWS * Q40T_TILE_ROWS;
    return d_pad / Q40T_TILE_ROWS * (long long)(n / Q40T_TILE_K) * Q40T_TILE_BYTES;
}

__global__ void q40_repack_kernel(const uint8_t* __restrict__ raw, int rows, int nb, uint8_t* __restrict__ out, int KT,
                                  int dst_row0, int dst_row_step) {
    // one thread per (source row, block); 18-byte source blocks are only 2-byte aligned
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)rows * nb) return;
    const int r = (int)(idx / nb), b = (int)(idx % nb);
    const uint16_t* src = reinterpret_cast<const uint16_t*>(raw + ((size_t)r * nb + b) * 18);
    uint16_t v[9];
#pragma unroll
    for (int i = 0; i < 9; i++) v[i] = src[i];
    const int drow = dst_row0 + r * dst_row_step;
    const int rt = drow >> 7, rr = drow & 127;
    const int kt = b >> 2, kb = b & 3;
    uint8_t* tile = out + ((size_t)rt * KT + kt) * Q40T_TILE_BYTES;
    uint4 q;
    q.x = (uint32_t)v[1] | ((uint32_t)v[2] << 16);
    q.y = (uint32_t)v[3] | ((uint32_t)v[4] << 16);
    q.z = (uint32_t)v[5] | ((uint32_t)v[6] << 16);
    q.w = (uint32_t)v[7] | ((uint32_t)v[8] << 16);
    *reinterpret_cast<uint4*>(tile + rr * 64 + (((kb + (rr >> 1)) & 3) << 4)) = q;
    *reinterpret_cast<uint16_t*>(tile + Q40T_SCALE_OFF + rr * 8 + kb * 2) = v[0];
}

extern "C" int dl_q40_repack(const uint8_t* raw, int rows, int n, uint8_t* q40t, int d_pad, int dst_row0,
                             int dst_row_step, void* stream) {
    if (n % Q40T_TILE_K != 0) { dl_set_error("dl_q40_repack: n=%d not a multiple of 128", n); return -1; }
    if (dst_row0 + (rows - 1) * dst_row_step >= d_pad) { dl_set_error("dl_q40_repack: rows exceed destination"); return -1; }
    const int nb = n / 32;
    const long long total = (long long)rows * nb;
    const int threads = 256;
    const unsigned blocks = (unsigned)((total + threads - 1) / threads);
    q40_repack_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(raw, rows, nb, q40t, n / Q40T_TILE_K, dst_row0,
                                                                    dst_row_step);
    CHECK_LAUNCH("dl_q40_repack");
    return 0;
}

__global__ void q40t_dequant_kernel(const uint8_t* __restrict__ w, int d, int n, float* __restrict__ out) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // one thread per (row, block)
    const int nb = n / 32;
    if (idx >= (long long)d * nb) return;
    const int r = (int)(idx / nb), b = (int)(idx % nb);
    const int rt = r >> 7, rr = r & 127, kt = b >> 2, kb = b & 3;
    const uint8_t* tile = w + ((size_t)rt * (n / Q40T_TILE_K) + kt) * Q40T_TILE_BYTES;
    const uint8_t* q = tile + rr * 64 + (((kb + (rr >> 1)) & 3) << 4);
    const float dw = __half2float(*reinterpret_cast<const __half*>(tile + Q40T_SCALE_OFF + rr * 8 + kb * 2));
    float* o = out + (size_t)r * n + (size_t)b * 32;
    for (int j = 0; j < 16; j++) {
        o[j] = (float)((int)(q[j] & 0xF) - 8) * dw;
        o[j + 16] = (float)((int)(q[j] >> 4) - 8) * dw;
    }
}

extern "C" int dl_q40t_dequant(const uint8_t* q40t, int d, int n, float* out, void* stream) {
    const long long total = (long long)d * (n / 32);
    q40t_dequant_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(q40t, d, n, out);
    CHECK_LAUNCH("dl_q40t_dequant");
    return 0;
}

// ------------------------------------------------------------------------------------------------
// embedding gather (+ per-forward step counter used as the epoch base of all peer flags)
// ------------------------------------------------------------------------------------------------
__global__ void embedding_kernel(const float* __restrict__ table, const int* __restrict__ tokens, float* __restrict__ x,
                                 int dim, uint32_t* step_counter, unsigned long long* tr) {
    trace_begin(tr);
    pdl_launch_dependents();
    pdl_wait();
    const int m = blockIdx.x;
    const float4* src = reinterpret_cast<const float4*>(table + (size_t)tokens[m] * dim);
    float4* dst = reinterpret_cast<float4*>(x + (size_t)m * dim);
    for (int i = threadIdx.x; i < dim / 4; i += blockDim.x) dst[i] = src[i];
    if (step_counter != nullptr && m == 0 && threadIdx.x == 0) *step_counter = *step_counter + 1u;
    trace_end(tr);
}

extern "C" int dl_embedding(const float* table, const int* tokens, float* x, int M, int dim, uint32_t* step_counter,
                            void* stream) {
    embedding_kernel<<<M, 256, 0, (cudaStream_t)stream>>>(table, tokens, x, dim, step_counter, dl_trace_next(TR_EMBED));
    CHECK_LAUNCH("dl_embedding");
    return 0;
}

// ------------------------------------------------------------------------------------------------
// reduce (merge-add of the peers' partial sums) + residual + RMSNorm + Q80 cast
// one CTA per activation row; dim <= 16384
// ------------------------------------------------------------------------------------------------
#define RNQ_THREADS 512
#define RNQ_MAX_PER_THREAD 32     // dim <= 512*32 = 16384

struct RnqArgs {
    const float* x_in;
    float* x_out;
    const float* partials;      // [n_src][mcap_p][dim] with src_stride elements between sources
    int n_src;
    long long src_stride;
    const uint32_t* flags;
    const uint32_t* step_counter;
    uint32_t flag_mul, flag_add;
    int ll;                     // partials are LL words {f32 bits, epoch}: poll the data, no flags
    const float* norm_w;
    float eps;
    int dim, M, mcap;
    uint8_t* xq_out;
    uint8_t* xh_out;
    float* y_out;
    unsigned long long* tr;
};

__global__ void __launch_bounds__(RNQ_THREADS) reduce_norm_quant_kernel(const __grid_constant__ RnqArgs a) {
    trace_begin(a.tr);
    __shared__ float s_part[RNQ_THREADS / 32];
    __shared__ float s_inv;
    pdl_launch_dependents();
    pdl_wait();
    const int m = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (a.flags != nullptr && a.n_src > 0) {
        if (tid < a.n_src) {
            const uint32_t epoch = ld_volatile_u32(a.step_counter) * a.flag_mul + a.flag_add;
            wait_flag_ge(a.flags + tid, epoch);
        }
        __syncthreads();
    }
    const int ll = (a.ll == 1) ? 1 : 0;
    const bool wire = a.ll == 2;      // partials are Q80 wire rows: dim int8 + dim/32 fp16 scales (see epilogue.cuh)
    const size_t rowb = (size_t)a.dim + (a.dim >> 4);
    const size_t src_rows = (size_t)a.src_stride / a.dim;
    const uint32_t ll_epoch = ll ? ld_volatile_u32(a.step_counter) * a.flag_mul + a.flag_add : 0u;
    // pass 1: merged residual (float4, coalesced) -> x_out, sum of squares
    constexpr int nwarps = RNQ_THREADS / 32;
    const size_t row = (size_t)m * a.dim;
    float ss = 0.0f;
    for (int i4 = tid; i4 < a.dim / 4; i4 += RNQ_THREADS) {
        float4 v = reinterpret_cast<const float4*>(a.x_in + row)[i4];
        for (int s = 0; s < a.n_src; s++) {
            float4 p;
            if (wire) {
                const uint8_t* wr = reinterpret_cast<const uint8_t*>(a.partials) + ((size_t)s * src_rows + m) * rowb;
                const int q4 = *reinterpret_cast<const int*>(wr + 4 * i4);
                const float d = __half2float(*reinterpret_cast<const __half*>(wr + a.dim + ((i4 >> 3) << 1)));
                p = make_float4((float)(int8_t)(q4 & 0xff) * d, (float)(int8_t)((q4 >> 8) & 0xff) * d,
                                (float)(int8_t)((q4 >> 16) & 0xff) * d, (float)(q4 >> 24) * d);
            } else {
                p = part_load4(a.partials, (size_t)s * a.src_stride + row + 4 * (size_t)i4, ll, ll_epoch);
            }
            v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
        }
        if (a.x_out != nullptr) reinterpret_cast<float4*>(a.x_out + row)[i4] = v;
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    ss = warp_sum(ss);
    if (lane == 0) s_part[warp] = ss;
    __syncthreads();
    if (warp == 0) {
        float t = (lane < nwarps) ? s_part[lane] : 0.0f;
        t = warp_sum(t);
        if (lane == 0) s_inv = 1.0f / sqrtf(t / (float)a.dim + a.eps);
    }
    __syncthreads();
    const float inv = (a.norm_w != nullptr) ? s_inv : 1.0f;
    // pass 2: one thread per 32-block (independent 16-byte loads, no shuffles)
    for (int b = tid; b < a.dim / 32; b += RNQ_THREADS) {
        const size_t e = row + (size_t)b * 32;
        float4 v[8];
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = reinterpret_cast<const float4*>(a.x_in + e)[j];
        for (int s = 0; s < a.n_src; s++) {
            if (wire) {
                const uint8_t* wr = reinterpret_cast<const uint8_t*>(a.partials) + ((size_t)s * src_rows + m) * rowb;
                const uint4 qa = *reinterpret_cast<const uint4*>(wr + b * 32), qb = *reinterpret_cast<const uint4*>(wr + b * 32 + 16);
                const float d = __half2float(*reinterpret_cast<const __half*>(wr + a.dim + (b << 1)));
                const uint32_t w8[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const int q4 = (int)w8[j];
                    v[j].x += (float)(int8_t)(q4 & 0xff) * d; v[j].y += (float)(int8_t)((q4 >> 8) & 0xff) * d;
                    v[j].z += (float)(int8_t)((q4 >> 16) & 0xff) * d; v[j].w += (float)(q4 >> 24) * d;
                }
                continue;
            }
            const size_t pe = (size_t)s * a.src_stride + e;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float4 t = part_load4(a.partials, pe + 4 * j, ll, ll_epoch);
                v[j].x += t.x; v[j].y += t.y; v[j].z += t.z; v[j].w += t.w;
            }
        }
        float amax = 0.0f;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (a.norm_w != nullptr) {
                const float4 w4 = reinterpret_cast<const float4*>(a.norm_w + (size_t)b * 32)[j];
                v[j].x = w4.x * (inv * v[j].x); v[j].y = w4.y * (inv * v[j].y);
                v[j].z = w4.z * (inv * v[j].z); v[j].w = w4.w * (inv * v[j].w);
            }
            amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v[j].x), fabsf(v[j].y)), fmaxf(fabsf(v[j].z), fabsf(v[j].w))));
        }
        if (a.y_out != nullptr) {
#pragma unroll
            for (int j = 0; j < 8; j++) reinterpret_cast<float4*>(a.y_out + e)[j] = v[j];
        }
        if (a.xq_out == nullptr && a.xh_out == nullptr) continue;
        const float d = amax / 127.0f;
        const float id = (d != 0.0f) ? (1.0f / d) : 0.0f;
        const float df = __half2float(__float2half_rn(d));
        int q[32];
        int sumq = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            q[4 * j + 0] = __float2int_rn(v[j].x * id); q[4 * j + 1] = __float2int_rn(v[j].y * id);
            q[4 * j + 2] = __float2int_rn(v[j].z * id); q[4 * j + 3] = __float2int_rn(v[j].w * id);
            sumq += q[4 * j] + q[4 * j + 1] + q[4 * j + 2] + q[4 * j + 3];
        }
        const int kt = b >> 2, kb = b & 3;
        if (a.xq_out != nullptr) {
            uint8_t* rowp = a.xq_out + ((size_t)kt * a.mcap + m) * XQ_ROW_BYTES;
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 8; j++)
                pk[j] = (uint32_t)(q[4 * j] & 0xff) | ((uint32_t)(q[4 * j + 1] & 0xff) << 8) |
                        ((uint32_t)(q[4 * j + 2] & 0xff) << 16) | ((uint32_t)(q[4 * j + 3] & 0xff) << 24);
            uint4* qd = reinterpret_cast<uint4*>(rowp + kb * 32);
            qd[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            qd[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            reinterpret_cast<float*>(rowp + 128)[kb] = df;
            reinterpret_cast<int*>(rowp + 144)[kb] = 8 * sumq;
        }
        if (a.xh_out != nullptr) {
            // chunk w holds elements [4w, 4w+2, 4w+1, 4w+3, 4w+16, 4w+18, 4w+17, 4w+19] (see xh_slot)
            uint8_t* base = a.xh_out + (size_t)kt * (size_t)(a.mcap >> 3) * XH_ROWGROUP_BYTES +
                            (size_t)(m >> 3) * XH_ROWGROUP_BYTES + (size_t)(kb * 4) * 128 + (size_t)(m & 7) * 16;
            const float dfh = fminf(df, 515.0f);       // fp16 operand: saturate (|x| > 65405) instead of producing inf
#pragma unroll
            for (int w = 0; w < 4; w++) {
                const int o = 4 * w;
                __half2 h0 = __floats2half2_rn((float)q[o] * dfh, (float)q[o + 2] * dfh);
                __half2 h1 = __floats2half2_rn((float)q[o + 1] * dfh, (float)q[o + 3] * dfh);
                __half2 h2 = __floats2half2_rn((float)q[o + 16] * dfh, (float)q[o + 18] * dfh);
                __half2 h3 = __floats2half2_rn((float)q[o + 17] * dfh, (float)q[o + 19] * dfh);
                uint4 u;
                u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
                u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
                *reinterpret_cast<uint4*>(base + w * 128) = u;
            }
        }
    }
    trace_end(a.tr);
}

extern "C" int dl_reduce_norm_quant(const float* x_in, float* x_out, const float* partials, int n_src, int src_stride,
                                    const uint32_t* flags, const uint32_t* step_counter, uint32_t flag_mul,
                                    uint32_t flag_add, int ll, const float* norm_w, float eps, int dim, int M, int mcap,
                                    uint8_t* xq_out, uint8_t* xh_out, float* y_out, int pdl, void* stream) {
    if (dim % 32 != 0 || dim > RNQ_THREADS * RNQ_MAX_PER_THREAD) { dl_set_error("dl_reduce_norm_quant: bad dim %d", dim); return -1; }
    if (n_src > DL_MAX_PEERS) { dl_set_error("dl_reduce_norm_quant: n_src > 8"); return -1; }
    RnqArgs a;
    a.x_in = x_in; a.x_out = x_out; a.partials = partials; a.n_src = n_src; a.src_stride = src_stride;
    a.flags = flags; a.step_counter = step_counter; a.flag_mul = flag_mul; a.flag_add = flag_add; a.ll = ll;
    a.norm_w = norm_w; a.eps = eps; a.dim = dim; a.M = M; a.mcap = mcap;
    a.xq_out = xq_out; a.xh_out = xh_out; a.y_out = y_out;
    a.tr = dl_trace_next(TR_RNQ);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(M);
    cfg.blockDim = dim3(RNQ_THREADS);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    if (pdl) {
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
    }
    cudaError_t e = cudaLaunchKernelEx(&cfg, reduce_norm_quant_kernel, a);
    if (e != cudaSuccess) { dl_set_error("dl_reduce_norm_quant: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}

extern "C" int dl_quant_q80(const float* x, int M, int K, int mcap, uint8_t* xq_out, uint8_t* xh_out, void* stream) {
    // the same kernel with no residual/partials/norm: y = x
    return dl_reduce_norm_quant(x, nullptr, nullptr, 0, 0, nullptr, nullptr, 0, 0, 0, nullptr, 0.0f, K, M, mcap, xq_out,
                                xh_out, nullptr, 0, stream);
}

// ------------------------------------------------------------------------------------------------
// greedy sampling on device (+ optional wait for the peers' logits slices, + device-resident
// decode loop support: feed the token back, advance the position, record a history)
// ------------------------------------------------------------------------------------------------
struct ArgmaxArgs {
    const float* logits;
    int vocab, ld;
    int* tokens_out;
    const uint32_t* flags;
    int n_flags;
    const uint32_t* step_counter;
    uint32_t flag_mul, flag_add;
    int* pos_inc;
    int* history;
    int history_cap;
    uint8_t* scratch;
    unsigned long long* tr;
};

#define ARGMAX_CHUNKS 32
#define ARGMAX_THREADS 256

DL_DEVINL void argmax_combine(float& best, int& bi, float ov, int oi) {
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
}

// grid (ARGMAX_CHUNKS, M): every CTA reduces one vocabulary chunk; the last CTA of a row (self-
// resetting counter in scratch) reduces the chunk winners and publishes the token.
__global__ void __launch_bounds__(ARGMAX_THREADS) argmax_kernel(const __grid_constant__ ArgmaxArgs a) {
    __shared__ float s_v[ARGMAX_THREADS / 32];
    __shared__ int s_i[ARGMAX_THREADS / 32];
    __shared__ int s_last;
    trace_begin(a.tr);
    pdl_launch_dependents();
    pdl_wait();
    if (a.flags != nullptr) {
        if ((int)threadIdx.x < a.n_flags) {
            const uint32_t epoch = ld_volatile_u32(a.step_counter) * a.flag_mul + a.flag_add;
            wait_flag_ge(a.flags + threadIdx.x, epoch);
        }
        __syncthreads();
    }
    const int m = blockIdx.y, ch = blockIdx.x, M = gridDim.y;
    unsigned* counters = reinterpret_cast<unsigned*>(a.scratch);                       // [M]
    float* part_v = reinterpret_cast<float*>(a.scratch + 1024);                        // [M][CHUNKS]
    int* part_i = reinterpret_cast<int*>(a.scratch + 1024 + (size_t)M * ARGMAX_CHUNKS * 4);
    const int per = ((a.vocab + ARGMAX_CHUNKS - 1) / ARGMAX_CHUNKS + 3) & ~3;
    const int i0 = ch * per, i1 = min(a.vocab, i0 + per);
    const float* row = a.logits + (size_t)m * a.ld;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = i0 + threadIdx.x; i < i1; i += ARGMAX_THREADS) argmax_combine(best, bi, row[i], i);
    for (int o = 16; o > 0; o >>= 1)
        argmax_combine(best, bi, __shfl_xor_sync(0xffffffffu, best, o), __shfl_xor_sync(0xffffffffu, bi, o));
    if ((threadIdx.x & 31) == 0) { s_v[threadIdx.x >> 5] = best; s_i[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < ARGMAX_THREADS / 32; w++) argmax_combine(best, bi, s_v[w], s_i[w]);
        part_v[m * ARGMAX_CHUNKS + ch] = best;
        part_i[m * ARGMAX_CHUNKS + ch] = bi;
        __threadfence();
        const unsigned prev = atomicAdd(&counters[m], 1u);
        s_last = (prev == ARGMAX_CHUNKS - 1) ? 1 : 0;
        if (s_last) counters[m] = 0u;
    }
    __syncthreads();
    if (!s_last) { trace_end(a.tr); return; }
    if (threadIdx.x < 32) {
        __threadfence();
        best = __ldcg(&part_v[m * ARGMAX_CHUNKS + threadIdx.x]);
        bi = __ldcg(&part_i[m * ARGMAX_CHUNKS + threadIdx.x]);
        for (int o = 16; o > 0; o >>= 1)
            argmax_combine(best, bi, __shfl_xor_sync(0xffffffffu, best, o), __shfl_xor_sync(0xffffffffu, bi, o));
        if (threadIdx.x == 0) {
            if (bi == 0x7fffffff) bi = 0;      // all-NaN row: stay in range
            a.tokens_out[m] = bi;
            if (a.pos_inc != nullptr) a.pos_inc[m] += 1;
            if (a.history != nullptr) {
                const uint32_t c = ld_volatile_u32(a.step_counter);
                a.history[(size_t)((c - 1u) % (uint32_t)a.history_cap) * M + m] = bi;
            }
        }
    }
    trace_end(a.tr);
}

extern "C" long long dl_argmax_scratch_bytes(int M) { return 1024 + (long long)M * ARGMAX_CHUNKS * 8; }

extern "C" int dl_argmax(const float* logits, int M, int vocab, int ld, int* tokens_out, const uint32_t* flags,
                         int n_flags, const uint32_t* step_counter, uint32_t flag_mul, uint32_t flag_add, int* pos_inc,
                         int* history, int history_cap, void* scratch, int pdl, void* stream) {
    static_assert(ARGMAX_CHUNKS == 32, "final reduce assumes one warp of chunk winners");
    ArgmaxArgs a;
    a.logits = logits; a.vocab = vocab; a.ld = ld; a.tokens_out = tokens_out;
    a.flags = flags; a.n_flags = n_flags; a.step_counter = step_counter; a.flag_mul = flag_mul; a.flag_add = flag_add;
    a.pos_inc = pos_inc; a.history = history; a.history_cap = history_cap;
    a.scratch = reinterpret_cast<uint8_t*>(scratch);
    a.tr = dl_trace_next(TR_ARGMAX);
    if (history != nullptr && step_counter == nullptr) { dl_set_error("dl_argmax: history needs step_counter"); return -1; }
    if (scratch == nullptr) { dl_set_error("dl_argmax: needs zero-initialised scratch (dl_argmax_scratch_bytes(min(M, 256)))"); return -1; }
    if (M > 256 && history != nullptr) { dl_set_error("dl_argmax: the device-loop history supports M <= 256"); return -1; }
    // the scratch holds 256 rows of counters and chunk winners: larger batches go through it 256 rows at a time
    // (launches on one stream run in order, and the counters return to zero)
    for (int m0 = 0; m0 < M; m0 += 256) {
        const int mb = (M - m0 < 256) ? M - m0 : 256;
        a.logits = logits + (size_t)m0 * ld;
        a.tokens_out = tokens_out + m0;
        a.pos_inc = pos_inc != nullptr ? pos_inc + m0 : nullptr;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(ARGMAX_CHUNKS, mb);
        cfg.blockDim = dim3(ARGMAX_THREADS);
        cfg.stream = (cudaStream_t)stream;
        cudaLaunchAttribute attr[1];
        if (pdl && m0 == 0) {
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
        }
        cudaError_t e = cudaLaunchKernelEx(&cfg, argmax_kernel, a);
        if (e != cudaSuccess) { dl_set_error("dl_argmax: %s", cudaGetErrorString(e)); return -(int)e; }
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------
// dl_argmax_noalloc: greedy pick for callers that must not allocate device memory.
//
// Who this is for: code that runs where cudaMalloc is off limits -- a plugin inside a host
// application that owns the whole memory budget, a latency-critical serving loop that forbids
// allocator calls after start-up, or a caller that only holds raw pointers and no allocator at
// all.  dl_argmax() needs a caller-provided scratch buffer (chunk winners + completion counters);
// this variant keeps that scratch in a fixed array inside the library instead, so the caller
// passes nothing but its logits and its output.
//
// Limits and rules:
//  * at most DL_ARGMAX_NOALLOC_MAX_ROWS (64) rows per call; larger M is refused with an error.
//    The array (17 KB) is part of the loaded module: every GPU that loads this library holds it
//    until the process exits.
//  * the first call on each device resolves the array's device address (cudaGetSymbolAddress, with
//    that device current).  Make that first call before starting any stream capture.
//  * all calls on a device share the one array: two calls whose kernels may overlap on the GPU
//    (different streams, no ordering between them) race on it.  Serialise them.
//  * the array keeps whatever the previous call left; the counters the kernel relies on being zero
//    are cleared on the caller's stream at the start of every call (cudaMemsetAsync, no allocation).
//  * no cudaMalloc / cudaFree / allocating library routine is called, here or in the kernel.
// ------------------------------------------------------------------------------------------------
#define DL_ARGMAX_NOALLOC_MAX_ROWS 64
#define DL_ARGMAX_NOALLOC_BYTES (1024 + DL_ARGMAX_NOALLOC_MAX_ROWS * ARGMAX_CHUNKS * 8)
static __device__ __align__(256) unsigned char g_argmax_noalloc_scratch[DL_ARGMAX_NOALLOC_BYTES];

extern "C" int dl_argmax_noalloc(const float* logits, int M, int vocab, int ld, int* tokens_out, void* stream) {
    // In host code `g_argmax_noalloc_scratch` is only a handle for the runtime, not a device address:
    // the real address is looked up once per device and cached here.
    static unsigned char* dev_addr[64] = {nullptr};
    if (M < 1 || M > DL_ARGMAX_NOALLOC_MAX_ROWS) {
        dl_set_error("dl_argmax_noalloc: M=%d is outside 1..%d (the built-in scratch is sized for %d rows; use "
                     "dl_argmax with your own scratch for more)", M, DL_ARGMAX_NOALLOC_MAX_ROWS, DL_ARGMAX_NOALLOC_MAX_ROWS);
        return -1;
    }
    if (logits == nullptr || tokens_out == nullptr || vocab < 1 || ld < vocab) {
        dl_set_error("dl_argmax_noalloc: bad arguments (logits/tokens_out NULL, vocab < 1 or ld < vocab)");
        return -1;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        dl_set_error("dl_argmax_noalloc: cudaGetDevice: %s", e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range");
        return e != cudaSuccess ? -(int)e : -1;
    }
    if (dev_addr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_argmax_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr) {
            dl_set_error("dl_argmax_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return e != cudaSuccess ? -(int)e : -1;
        }
        dev_addr[dev] = reinterpret_cast<unsigned char*>(p);
    }
    unsigned char* scratch = dev_addr[dev];
    // completion counters (first 1 KB) must be zero when the kernel starts
    e = cudaMemsetAsync(scratch, 0, 1024, (cudaStream_t)stream);
    if (e != cudaSuccess) { dl_set_error("dl_argmax_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(e)); return -(int)e; }
    ArgmaxArgs a;
    a.logits = logits; a.vocab = vocab; a.ld = ld; a.tokens_out = tokens_out;
    a.flags = nullptr; a.n_flags = 0; a.step_counter = nullptr; a.flag_mul = 0; a.flag_add = 0;
    a.pos_inc = nullptr; a.history = nullptr; a.history_cap = 0;
    a.scratch = scratch;
    a.tr = nullptr;
    argmax_kernel<<<dim3(ARGMAX_CHUNKS, M), ARGMAX_THREADS, 0, (cudaStream_t)stream>>>(a);
    e = cudaGetLastError();
    if (e != cudaSuccess) { dl_set_error("dl_argmax_noalloc: launch: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}

// wait until flags[i] >= epoch for i < n (orders later stream work after the peers' pushes)
__global__ void wait_flags_kernel(const uint32_t* flags, int n, const uint32_t* step_counter, uint32_t mul, uint32_t add) {
    pdl_launch_dependents();
    pdl_wait();
    if ((int)threadIdx.x < n) {
        const uint32_t epoch = ld_volatile_u32(step_counter) * mul + add;
        wait_flag_ge(flags + threadIdx.x, epoch);
    }
}

extern "C" int dl_wait_flags(const uint32_t* flags, int n, const uint32_t* step_counter, uint32_t flag_mul,
                             uint32_t flag_add, void* stream) {
    wait_flags_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(flags, n, step_counter, flag_mul, flag_add);
    CHECK_LAUNCH("dl_wait_flags");
    return 0;
}
