// The elementwise / data-movement kernels of the f32-weights path (`--buffer-float-type f32`: no Q80 cast anywhere).
// Together with gemm_f32_kernel (matmul), reduce_norm_quant_kernel (merge-add + RMSNorm, f32 output), attention_kernel
// (paged f32 KV cache), embedding_kernel and argmax_kernel they make that path native: no cuBLAS, no NCCL, no eager
// PyTorch ops between the embedding and the logits.
//   rope_kv_f32    OP_ROPE_LLAMA + OP_SHIFT   (ropeLlamaForward_F32_F32, shiftForward_F32_F32: nn-cpu-ops.cpp:1090-1120,1253-1275)
//   silu_mul_f32   OP_SILU / OP_GELU + OP_MUL (nn-cpu-ops.cpp:1058-1078,1163-1183)
//   push_f32       SYNC_NODE_SLICES of f32 slices (nn-network.cpp:537-569): this rank's partial sums (or logits slice)
//                  stored into every rank's exchange buffer + release flags; the merge-add (mergeAddForward_F32_F32,
//                  nn-cpu-ops.cpp:835-852) happens in the consumer (reduce_norm_quant_kernel), in rank order.
#include "epilogue.cuh"

extern void dl_set_error(const char* fmt, ...);

#define F32_CHECK(name)                                                                  \
    do {                                                                                 \
        cudaError_t e__ = cudaGetLastError();                                            \
        if (e__ != cudaSuccess) { dl_set_error("%s: %s", name, cudaGetErrorString(e__)); return -(int)e__; } \
    } while (0)

struct RopeKvArgs {
    const float* qkv;          // [M][q_dim + 2 kv_dim]
    float* q_out;              // [M][q_dim]
    void* k_cache;             // [pages][n_kv][32][hs]
    void* v_cache;
    const float* freqs;        // [hs/2]
    const int* pos;
    const int* slot;
    const int* block_table;
    int M, q_dim, kv_dim, hs, n_kv, max_pages, kv_f32;
};

// one thread per (row, pair of adjacent dims)
__global__ void rope_kv_f32_kernel(const __grid_constant__ RopeKvArgs a) {
    const int d = a.q_dim + 2 * a.kv_dim;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)a.M * (d >> 1)) return;
    const int m = (int)(idx / (d >> 1)), pr = (int)(idx - (long long)m * (d >> 1));
    const int row = 2 * pr;
    const float2 v = *reinterpret_cast<const float2*>(a.qkv + (size_t)m * d + row);
    float v0 = v.x, v1 = v.y;
    const int region = (row < a.q_dim) ? 0 : ((row < a.q_dim + a.kv_dim) ? 1 : 2);
    const int rrow = (region == 0) ? row : ((region == 1) ? row - a.q_dim : row - a.q_dim - a.kv_dim);
    const int hd = rrow % a.hs, kvh = rrow / a.hs;
    const int p = a.pos[m];
    if (region < 2) {
        float sn, cs;
        sincosf((float)p * a.freqs[hd >> 1], &sn, &cs);
        const float r0 = v0 * cs - v1 * sn, r1 = v0 * sn + v1 * cs;
        v0 = r0; v1 = r1;
    }
    if (region == 0) {
        *reinterpret_cast<float2*>(a.q_out + (size_t)m * a.q_dim + rrow) = make_float2(v0, v1);
    } else {
        const int page = a.block_table[(size_t)a.slot[m] * a.max_pages + p / 32];
        const size_t off = (((size_t)page * a.n_kv + kvh) * 32 + (p % 32)) * a.hs + hd;
        void* base = (region == 1) ? a.k_cache : a.v_cache;
        if (a.kv_f32) *reinterpret_cast<float2*>(reinterpret_cast<float*>(base) + off) = make_float2(v0, v1);
        else *reinterpret_cast<__half2*>(reinterpret_cast<__half*>(base) + off) = __floats2half2_rn(v0, v1);
    }
}

extern "C" int dl_f32_rope_kv(const float* qkv, int M, int q_dim, int kv_dim, int head_size, int n_kv_heads, const float* freqs,
                              const int* pos, const int* slot, const int* block_table, int max_pages, float* q_out,
                              void* k_cache, void* v_cache, int kv_f32, void* stream) {
    if ((q_dim | kv_dim | head_size) & 1) { dl_set_error("dl_f32_rope_kv: odd dimensions"); return -1; }
    RopeKvArgs a;
    a.qkv = qkv; a.q_out = q_out; a.k_cache = k_cache; a.v_cache = v_cache; a.freqs = freqs; a.pos = pos; a.slot = slot;
    a.block_table = block_table; a.M = M; a.q_dim = q_dim; a.kv_dim = kv_dim; a.hs = head_size; a.n_kv = n_kv_heads;
    a.max_pages = max_pages; a.kv_f32 = kv_f32;
    const long long total = (long long)M * ((q_dim + 2 * kv_dim) >> 1);
    rope_kv_f32_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(a);
    F32_CHECK("dl_f32_rope_kv");
    return 0;
}

// out[m][j] = act(g[m][j]) * u[m][j]
__global__ void silu_mul_f32_kernel(const float* __restrict__ g, const float* __restrict__ u, float* __restrict__ out, long long n,
                                    int gelu) {
    const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    const float4 a = *reinterpret_cast<const float4*>(g + i), b = *reinterpret_cast<const float4*>(u + i);
    float4 o;
    o.x = (gelu ? act_gelu(a.x) : act_silu(a.x)) * b.x;
    o.y = (gelu ? act_gelu(a.y) : act_silu(a.y)) * b.y;
    o.z = (gelu ? act_gelu(a.z) : act_silu(a.z)) * b.z;
    o.w = (gelu ? act_gelu(a.w) : act_silu(a.w)) * b.w;
    *reinterpret_cast<float4*>(out + i) = o;
}

extern "C" int dl_f32_silu_mul(const float* g, const float* u, float* out, long long n, int gelu, void* stream) {
    if (n % 4 != 0) { dl_set_error("dl_f32_silu_mul: n must be a multiple of 4"); return -1; }
    silu_mul_f32_kernel<<<(unsigned)((n / 4 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(g, u, out, n, gelu);
    F32_CHECK("dl_f32_silu_mul");
    return 0;
}

struct PushArgs {
    const float* src;          // [M][w], row stride ld_src
    float* dst[DL_MAX_PEERS];  // destination buffers (peer-mapped), element offset dst_off, row stride ld_dst
    uint32_t* flag[DL_MAX_PEERS];
    int n_dst, M, w, ld_src, ld_dst;
    long long dst_off;
    const uint32_t* step_counter;
    uint32_t flag_mul, flag_add;
    uint32_t* cta_counter;
};

// grid = M CTAs (one row each).  Every CTA fences its peer stores; the last one to finish releases the flags.
__global__ void push_f32_kernel(const __grid_constant__ PushArgs a) {
    __shared__ int s_last;
    const int m = blockIdx.x;
    const float4* s = reinterpret_cast<const float4*>(a.src + (size_t)m * a.ld_src);
    for (int p = 0; p < a.n_dst; p++) {
        float4* d = reinterpret_cast<float4*>(a.dst[p] + a.dst_off + (size_t)m * a.ld_dst);
        for (int i = threadIdx.x; i < a.w / 4; i += blockDim.x) d[i] = s[i];
    }
    if (a.flag[0] == nullptr) return;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned prev = atomicAdd(a.cta_counter, 1u);
        s_last = (prev == (unsigned)a.M - 1u) ? 1 : 0;
        if (s_last) {
            __threadfence_system();
            *a.cta_counter = 0u;
        }
    }
    __syncthreads();
    if (s_last && (int)threadIdx.x < a.n_dst) {
        const uint32_t epoch = ld_volatile_u32(a.step_counter) * a.flag_mul + a.flag_add;
        st_release_sys(a.flag[threadIdx.x], epoch);
    }
}

extern "C" int dl_f32_push(const float* src, int M, int w, int ld_src, float* const* dst /*host array*/, int n_dst,
                           long long dst_off, int ld_dst, uint32_t* const* flags /*host array or NULL*/,
                           const uint32_t* step_counter, uint32_t flag_mul, uint32_t flag_add, uint32_t* cta_counter,
                           void* stream) {
    if (n_dst < 1 || n_dst > DL_MAX_PEERS || w % 4 != 0 || ld_src % 4 != 0 || ld_dst % 4 != 0 || dst_off % 4 != 0) {
        dl_set_error("dl_f32_push: bad arguments (n_dst 1..8, widths / strides / offset multiples of 4)");
        return -1;
    }
    PushArgs a;
    for (int p = 0; p < DL_MAX_PEERS; p++) {
        a.dst[p] = p < n_dst ? dst[p] : nullptr;
        a.flag[p] = (flags != nullptr && p < n_dst) ? flags[p] : nullptr;
    }
    a.src = src; a.n_dst = n_dst; a.M = M; a.w = w; a.ld_src = ld_src; a.ld_dst = ld_dst; a.dst_off = dst_off;
    a.step_counter = step_counter; a.flag_mul = flag_mul; a.flag_add = flag_add; a.cta_counter = cta_counter;
    push_f32_kernel<<<M, 256, 0, (cudaStream_t)stream>>>(a);
    F32_CHECK("dl_f32_push");
    return 0;
}
