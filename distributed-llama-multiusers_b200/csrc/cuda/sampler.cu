// Per-row token sampling on the device: temperature -> softmax -> top-p (nucleus) or multinomial,
// argmax at temperature 0 (reference: Sampler::sample, src/tokenizer.cpp:382-510, run there on the
// host over logits it already has in RAM).  With 32 users the host path means a 16 MB device-to-host
// copy and 32 sorts of a 128k-entry vocabulary per step; here only the token ids leave the GPU.
//
// One CTA per row.  The random coin of every row comes from the request's own host-side xorshift*
// generator (same stream of coins as the reference would consume), so a request stays reproducible
// for a given seed regardless of what else is in the batch.  Differences from the host sampler, both
// distribution-preserving: the nucleus is found by a 3-pass radix select on the probability bits
// instead of a sort (ties at the threshold are all kept), and the token is picked in index order
// inside the nucleus instead of descending-probability order.  All sums are 64-bit fixed point
// (2^-40 resolution), so the result does not depend on the order in which threads add.
#include "common.cuh"

#define SMP_THREADS 1024
#define SMP_BINS 2048
#define SMP_FIX 1099511627776.0f      // 2^40

struct SampleArgs {
    const float* logits;
    int M, vocab, ld;
    const float* temperature;
    const float* topp;
    const float* coin;
    int* tokens;
};

DL_DEVINL unsigned long long to_fix(float p) { return (unsigned long long)(p * SMP_FIX); }

// block-wide sum of a 64-bit value (deterministic: integer adds)
DL_DEVINL unsigned long long block_sum_u64(unsigned long long v, unsigned long long* red, int tid) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    unsigned long long s = 0;
    for (int w = 0; w < SMP_THREADS / 32; w++) s += red[w];
    return s;
}

__global__ void __launch_bounds__(SMP_THREADS) sample_kernel(const SampleArgs a) {
    __shared__ unsigned long long hist[SMP_BINS];
    __shared__ unsigned long long red[SMP_THREADS / 32];
    __shared__ unsigned long long chunk_sum[SMP_THREADS];
    __shared__ float redf[SMP_THREADS / 32];
    __shared__ int redi[SMP_THREADS / 32];
    __shared__ unsigned int sel_bin;
    __shared__ unsigned long long sel_above;
    __shared__ int sel_chunk;
    __shared__ unsigned long long sel_rest;
    const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* x = a.logits + (size_t)row * a.ld;
    const int V = a.vocab;
    const float temp = a.temperature[row];

    // ---- max (and its first index: the greedy pick) ----
    float m = -INFINITY;
    int mi = 0x7fffffff;
    for (int i = tid; i < V; i += SMP_THREADS) {
        const float v = x[i];
        if (v > m) { m = v; mi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, m, o);
        const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
        if (om > m || (om == m && oi < mi)) { m = om; mi = oi; }
    }
    if (lane == 0) { redf[warp] = m; redi[warp] = mi; }
    __syncthreads();
    m = redf[0]; mi = redi[0];
    for (int w = 1; w < SMP_THREADS / 32; w++)
        if (redf[w] > m || (redf[w] == m && redi[w] < mi)) { m = redf[w]; mi = redi[w]; }
    if (temp == 0.0f) {
        if (tid == 0) a.tokens[row] = mi;
        return;
    }
    const float inv_t = 1.0f / temp;
    auto prob = [&](int i) { return expf((x[i] - m) * inv_t); };      // unnormalised, in (0, 1]

    // ---- Z = sum of probabilities ----
    unsigned long long zp = 0;
    for (int i = tid; i < V; i += SMP_THREADS) zp += to_fix(prob(i));
    const unsigned long long Z = block_sum_u64(zp, red, tid);

    // ---- nucleus threshold (bit pattern of the smallest probability kept) ----
    const float topp = a.topp[row];
    unsigned int thr = 0;                       // keep everything
    if (topp > 0.0f && topp < 1.0f) {
        const unsigned long long target = (unsigned long long)((double)Z * (double)topp);
        unsigned int prefix = 0, prefix_mask = 0;
        unsigned long long above = 0;           // mass of the probabilities strictly above the current prefix range
        const int shifts[3] = {21, 10, 0};
        const int widths[3] = {11, 11, 10};
        for (int pass = 0; pass < 3; pass++) {
            const int sh = shifts[pass];
            const unsigned int bmask = (1u << widths[pass]) - 1u;
            for (int b = tid; b < SMP_BINS; b += SMP_THREADS) hist[b] = 0;
            __syncthreads();
            for (int i = tid; i < V; i += SMP_THREADS) {
                const float p = prob(i);
                const unsigned int bits = __float_as_uint(p);
                if ((bits & prefix_mask) == prefix) atomicAdd(&hist[(bits >> sh) & bmask], to_fix(p));
            }
            __syncthreads();
            if (tid == 0) {
                unsigned long long cum = above;
                int b = (int)bmask;
                for (; b > 0; b--) {            // from the largest probabilities down
                    if (cum + hist[b] >= target) break;
                    cum += hist[b];
                }
                sel_bin = (unsigned int)b;
                sel_above = cum;
            }
            __syncthreads();
            prefix |= sel_bin << sh;
            prefix_mask |= bmask << sh;
            above = sel_above;
            __syncthreads();
        }
        thr = prefix;
    }

    // ---- draw: index-ordered cumulative over the kept tokens; thread t owns a contiguous index chunk ----
    const int C = (V + SMP_THREADS - 1) / SMP_THREADS;
    const int i0 = tid * C, i1 = min(V, i0 + C);
    unsigned long long mine = 0;
    for (int i = i0; i < i1; i++) {
        const float p = prob(i);
        if (__float_as_uint(p) >= thr) mine += to_fix(p);
    }
    chunk_sum[tid] = mine;
    __syncthreads();
    if (tid == 0) {
        unsigned long long total = 0;
        for (int t = 0; t < SMP_THREADS; t++) total += chunk_sum[t];
        unsigned long long want = (unsigned long long)((double)total * (double)a.coin[row]);
        if (want >= total) want = total ? total - 1 : 0;
        int t = 0;
        for (; t < SMP_THREADS - 1; t++) {
            if (want < chunk_sum[t]) break;
            want -= chunk_sum[t];
        }
        sel_chunk = t;
        sel_rest = want;
    }
    __syncthreads();
    if (tid == sel_chunk) {
        unsigned long long want = sel_rest;
        int pick = mi;                          // fallback (empty chunk through rounding): the most likely token
        for (int i = i0; i < i1; i++) {
            const float p = prob(i);
            if (__float_as_uint(p) < thr) continue;
            const unsigned long long f = to_fix(p);
            pick = i;
            if (want < f) break;
            want -= f;
        }
        a.tokens[row] = pick;
    }
}

extern void dl_set_error(const char* fmt, ...);

extern "C" int dl_sample_rows(const float* logits, int M, int vocab, int ld, const float* temperature, const float* topp,
                              const float* coin, int* tokens_out, void* stream) {
    if (M < 1 || vocab < 1) { dl_set_error("dl_sample_rows: empty problem"); return -1; }
    SampleArgs a{logits, M, vocab, ld, temperature, topp, coin, tokens_out};
    sample_kernel<<<M, SMP_THREADS, 0, (cudaStream_t)stream>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { dl_set_error("sample launch: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}
