// CPU backend: the whole Llama forward pass on host cores, tensor-parallel over TCP.
//
// This is the no-GPU configuration of the framework (BASELINE.json: "root + 1 worker over localhost TCP,
// q80 buffer, CPU (plumbing, no GPU)") and the native counterpart of the reference's CPU device, op kernels,
// socket layer, node synchroniser, config broadcast and weight distribution:
//
//   reference component                                         here
//   ---------------------------------------------------------   -------------------------------------------
//   quant codecs            src/nn/nn-quants.cpp:67-246         quantize_q80 / q40 unpack (scalar + AVX2)
//   op kernels              src/nn/nn-cpu-ops.cpp               embedding, rmsnorm, cast, matmul, rope, shift,
//                                                               multihead attention, silu/gelu, mul, merge-add
//   tinyBLAS Q40xQ80        src/nn/llamafile/sgemm.cpp:593-783  matmul_q40_q80 (weights unpacked once per block,
//                                                               reused for every batch row)
//   executor threads        src/nn/nn-executor.cpp:134-187      persistent pool (no thread spawn per forward)
//   slicers / splitters     src/nn/nn-core.cpp:198-303          slice_rows / slice_cols on Q40 blocks
//   socket layer + mesh     src/nn/nn-network.cpp:24-348        Mesh (blocking connect/accept, poll-driven exchange)
//   node synchroniser       src/nn/nn-network.cpp:510-601       broadcast_x (S1), all_gather (S2/S3), gather_logits (S4)
//   config broadcast        src/nn/nn-network.cpp:617-764       NodeConfig record (workers need no model file)
//   weight distribution     src/nn/nn-network.cpp:766-901       tagged slice records streamed to the owning worker
//   header / weight walker  src/llm.cpp:26-98,447-483           load_header / tensor offsets
//   mmap wrapper            src/mmap.hpp:35-81                  MappedFile
//   root / worker drivers   src/app.cpp:169-231,405-464         dl_cpu_forward / dl_cpu_worker_serve
//
// Independent design, not a translation: rows of a batch carry their own (position, KV slot) so that many users
// share a forward pass (the fork's D2/D3 defects cannot occur); a batch is synchronised in one exchange instead of
// row by row; sockets are driven by poll() so simultaneous sends cannot deadlock; the thread pool is persistent.
// Numerics follow SURVEY Appendix A/B (same quantisation points as models/oracle.py, which is the test oracle).
#include <arpa/inet.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/mman.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#if defined(__x86_64__)
#include <immintrin.h>
#define DL_X86 1
#endif

namespace {

constexpr int QK = 32;
constexpr int Q40_BYTES = 18, Q80_BYTES = 34;
enum FloatType { F_32 = 0, F_16 = 1, F_Q40 = 2, F_Q80 = 3 };

// ---- fp16 <-> fp32 (software; reference: nn-quants.cpp:12-65) ---------------------------------------
inline float f16_to_f32(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000) << 16;
    uint32_t exp = (h >> 10) & 0x1f, man = h & 0x3ff, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {
            int e = -1;
            do { man <<= 1; e++; } while (!(man & 0x400));
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ff) << 13);
        }
    } else if (exp == 31) bits = sign | 0x7f800000u | (man << 13);
    else bits = sign | ((exp + 112) << 23) | (man << 13);
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

inline uint16_t f32_to_f16(float f) {               // round to nearest even
    uint32_t x;
    memcpy(&x, &f, 4);
    const uint16_t sign = (uint16_t)((x >> 16) & 0x8000);
    x &= 0x7fffffffu;
    if (x > 0x7f800000u) return sign | 0x7e00;      // NaN
    if (x >= 0x47800000u) return sign | 0x7c00;     // >= 65536 (incl. inf)
    const uint32_t exp = x >> 23;
    uint32_t man = x & 0x7fffffu;
    if (exp < 113) {                                // below the smallest normal half (2^-14): subnormal or zero
        if (exp < 102) return sign;
        man |= 0x800000u;
        const int shift = 126 - (int)exp;           // 14..24: value = man * 2^(exp-150), half unit = 2^-24
        uint32_t h = man >> shift;
        const uint32_t rem = man & ((1u << shift) - 1), mid = 1u << (shift - 1);
        if (rem > mid || (rem == mid && (h & 1))) h++;
        return sign | (uint16_t)h;
    }
    uint32_t h = ((exp - 112) << 10) | (man >> 13);
    const uint32_t rem = man & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) h++;     // a carry into the exponent is the correct rounding
    return sign | (uint16_t)h;
}

// ---- Q80 quantise (reference: quantizeF32toQ80, nn-quants.cpp:67-173; half-to-even like its AVX2 path) ----
void quantize_q80_scalar(const float* x, uint8_t* out, int n) {
    for (int b = 0; b < n / QK; b++) {
        const float* g = x + b * QK;
        float amax = 0.f;
        for (int j = 0; j < QK; j++) amax = std::max(amax, fabsf(g[j]));
        const float d = amax / 127.f;
        const float id = d != 0.f ? 1.f / d : 0.f;
        uint8_t* o = out + (size_t)b * Q80_BYTES;
        const uint16_t dh = f32_to_f16(d);
        memcpy(o, &dh, 2);
        for (int j = 0; j < QK; j++) ((int8_t*)o)[2 + j] = (int8_t)nearbyintf(g[j] * id);
    }
}

void dequantize_q80(const uint8_t* in, float* y, int n) {
    for (int b = 0; b < n / QK; b++) {
        const uint8_t* p = in + (size_t)b * Q80_BYTES;
        uint16_t dh;
        memcpy(&dh, p, 2);
        const float d = f16_to_f32(dh);
        for (int j = 0; j < QK; j++) y[b * QK + j] = (float)((const int8_t*)p)[2 + j] * d;
    }
}

// y[m][r] = sum_blocks (sum_j (nib-8) * q) * d_w * d_x   (reference: matmul_Q80_Q40_F32, nn-cpu-ops.cpp:222-440)
void matmul_q40_q80_scalar(const uint8_t* w, int r0, int r1, int nb, const uint8_t* xq, int M, float* y, size_t ldy) {
    for (int r = r0; r < r1; r++) {
        const uint8_t* wr = w + (size_t)r * nb * Q40_BYTES;
        for (int m = 0; m < M; m++) {
            const uint8_t* xr = xq + (size_t)m * nb * Q80_BYTES;
            float acc = 0.f;
            for (int b = 0; b < nb; b++) {
                const uint8_t* wb = wr + (size_t)b * Q40_BYTES;
                const uint8_t* xb = xr + (size_t)b * Q80_BYTES;
                uint16_t dw, dx;
                memcpy(&dw, wb, 2);
                memcpy(&dx, xb, 2);
                const int8_t* q = (const int8_t*)xb + 2;
                int s = 0;
                for (int j = 0; j < 16; j++) {
                    s += ((int)(wb[2 + j] & 0x0f) - 8) * q[j];
                    s += ((int)(wb[2 + j] >> 4) - 8) * q[j + 16];
                }
                acc += (float)s * (f16_to_f32(dw) * f16_to_f32(dx));
            }
            y[(size_t)m * ldy + r] = acc;
        }
    }
}

#ifdef DL_X86
__attribute__((target("avx2,fma")))
void quantize_q80_avx2(const float* x, uint8_t* out, int n) {
    const __m256 sign_mask = _mm256_set1_ps(-0.0f);
    for (int b = 0; b < n / QK; b++) {
        const float* g = x + b * QK;
        __m256 v0 = _mm256_loadu_ps(g), v1 = _mm256_loadu_ps(g + 8), v2 = _mm256_loadu_ps(g + 16), v3 = _mm256_loadu_ps(g + 24);
        __m256 mx = _mm256_max_ps(_mm256_max_ps(_mm256_andnot_ps(sign_mask, v0), _mm256_andnot_ps(sign_mask, v1)),
                                  _mm256_max_ps(_mm256_andnot_ps(sign_mask, v2), _mm256_andnot_ps(sign_mask, v3)));
        __m128 m4 = _mm_max_ps(_mm256_extractf128_ps(mx, 1), _mm256_castps256_ps128(mx));
        m4 = _mm_max_ps(m4, _mm_movehl_ps(m4, m4));
        m4 = _mm_max_ss(m4, _mm_movehdup_ps(m4));
        const float amax = _mm_cvtss_f32(m4);
        const float d = amax / 127.f;
        const float id = d != 0.f ? 1.f / d : 0.f;
        uint8_t* o = out + (size_t)b * Q80_BYTES;
        const uint16_t dh = f32_to_f16(d);
        memcpy(o, &dh, 2);
        const __m256 mul = _mm256_set1_ps(id);
        const int rnd = _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC;
        __m256i i0 = _mm256_cvtps_epi32(_mm256_round_ps(_mm256_mul_ps(v0, mul), rnd));
        __m256i i1 = _mm256_cvtps_epi32(_mm256_round_ps(_mm256_mul_ps(v1, mul), rnd));
        __m256i i2 = _mm256_cvtps_epi32(_mm256_round_ps(_mm256_mul_ps(v2, mul), rnd));
        __m256i i3 = _mm256_cvtps_epi32(_mm256_round_ps(_mm256_mul_ps(v3, mul), rnd));
        i0 = _mm256_packs_epi32(i0, i1);
        i2 = _mm256_packs_epi32(i2, i3);
        i0 = _mm256_packs_epi16(i0, i2);
        i0 = _mm256_permutevar8x32_epi32(i0, _mm256_setr_epi32(0, 4, 1, 5, 2, 6, 3, 7));
        _mm256_storeu_si256((__m256i*)(o + 2), i0);
    }
}

// horizontal sum of 8 f32 lanes (macro: helpers called from target("avx2") functions would need the attribute too)
#define DL_HSUM8(dst, v) do { \
    __m128 s_ = _mm_add_ps(_mm256_extractf128_ps(v, 1), _mm256_castps256_ps128(v)); \
    s_ = _mm_add_ps(s_, _mm_movehl_ps(s_, s_)); \
    s_ = _mm_add_ss(s_, _mm_movehdup_ps(s_)); \
    (dst) = _mm_cvtss_f32(s_); } while (0)
// Q40 block -> 32 signed bytes (low nibbles = elements 0..15) and their absolute values
#define DL_UNPACK(wb, wv, aw) do { \
    const __m128i raw_ = _mm_loadu_si128((const __m128i*)((wb) + 2)); \
    wv = _mm256_set_m128i(_mm_srli_epi16(raw_, 4), raw_); \
    wv = _mm256_sub_epi8(_mm256_and_si256(wv, low), eight); \
    aw = _mm256_sign_epi8(wv, wv); } while (0)

// integer dot of one Q40 block (already unpacked to int8 `wv`, |wv| in `aw`) with one Q80 block, as 8 f32 lanes
#define DL_BLOCK_DOT(aw, wv, xptr) \
    _mm256_cvtepi32_ps(_mm256_madd_epi16(_mm256_maddubs_epi16(aw, _mm256_sign_epi8(_mm256_loadu_si256((const __m256i*)(xptr)), wv)), ones))

__attribute__((target("avx2,fma,f16c")))
void matmul_q40_q80_avx2(const uint8_t* w, int r0, int r1, int nb, const uint8_t* xq, int M, float* y, size_t ldy) {
    const __m256i low = _mm256_set1_epi8(0x0f), eight = _mm256_set1_epi8(8), ones = _mm256_set1_epi16(1);
    // activation scales widened once per call (they are reused by every weight row)
    std::vector<float> xd((size_t)M * nb);
    for (int m = 0; m < M; m++)
        for (int b = 0; b < nb; b++) {
            uint16_t h;
            memcpy(&h, xq + ((size_t)m * nb + b) * Q80_BYTES, 2);
            xd[(size_t)m * nb + b] = _cvtsh_ss(h);
        }
    if (M == 1) {                               // decode: two independent accumulator chains per row
        const float* sx = xd.data();
        for (int r = r0; r < r1; r++) {
            const uint8_t* wr = w + (size_t)r * nb * Q40_BYTES;
            __m256 a0 = _mm256_setzero_ps(), a1 = _mm256_setzero_ps();
            int b = 0;
            for (; b + 2 <= nb; b += 2) {
                const uint8_t* w0 = wr + (size_t)b * Q40_BYTES;
                const uint8_t* w1 = w0 + Q40_BYTES;
                uint16_t h0, h1;
                memcpy(&h0, w0, 2);
                memcpy(&h1, w1, 2);
                __m256i v0, u0, v1, u1;
                DL_UNPACK(w0, v0, u0);
                DL_UNPACK(w1, v1, u1);
                a0 = _mm256_fmadd_ps(_mm256_set1_ps(_cvtsh_ss(h0) * sx[b]), DL_BLOCK_DOT(u0, v0, xq + (size_t)b * Q80_BYTES + 2), a0);
                a1 = _mm256_fmadd_ps(_mm256_set1_ps(_cvtsh_ss(h1) * sx[b + 1]), DL_BLOCK_DOT(u1, v1, xq + (size_t)(b + 1) * Q80_BYTES + 2), a1);
            }
            for (; b < nb; b++) {
                const uint8_t* w0 = wr + (size_t)b * Q40_BYTES;
                uint16_t h0;
                memcpy(&h0, w0, 2);
                __m256i v0, u0;
                DL_UNPACK(w0, v0, u0);
                a0 = _mm256_fmadd_ps(_mm256_set1_ps(_cvtsh_ss(h0) * sx[b]), DL_BLOCK_DOT(u0, v0, xq + (size_t)b * Q80_BYTES + 2), a0);
            }
            a0 = _mm256_add_ps(a0, a1);
            DL_HSUM8(y[r], a0);
        }
        return;
    }
    // batch rows: weights unpacked once per block and reused for 4 rows.  Same arithmetic per row as the decode path
    // above (even blocks -> chain 0, odd blocks -> chain 1, chains added, then the horizontal sum), so a row's result
    // does not depend on which other rows share its batch: continuous batching stays bit-reproducible.
    constexpr int MB = 4;
    for (int r = r0; r < r1; r++) {
        const uint8_t* wr = w + (size_t)r * nb * Q40_BYTES;
        for (int m0 = 0; m0 < M; m0 += MB) {
            const int mc = std::min(MB, M - m0);
            __m256 acc0[MB], acc1[MB];
            for (int i = 0; i < MB; i++) { acc0[i] = _mm256_setzero_ps(); acc1[i] = _mm256_setzero_ps(); }
            for (int b = 0; b < nb; b++) {
                const uint8_t* wb = wr + (size_t)b * Q40_BYTES;
                uint16_t dwh;
                memcpy(&dwh, wb, 2);
                const float dw = _cvtsh_ss(dwh);
                __m256i wv, aw;
                DL_UNPACK(wb, wv, aw);
                __m256* acc = (b & 1) && b + 1 <= (nb & ~1) ? acc1 : acc0;      // a trailing odd block joins chain 0, as above
                for (int i = 0; i < mc; i++) {
                    const size_t xi = (size_t)(m0 + i) * nb + b;
                    acc[i] = _mm256_fmadd_ps(_mm256_set1_ps(dw * xd[xi]), DL_BLOCK_DOT(aw, wv, xq + xi * Q80_BYTES + 2), acc[i]);
                }
            }
            for (int i = 0; i < mc; i++) {
                const __m256 t = _mm256_add_ps(acc0[i], acc1[i]);
                DL_HSUM8(y[(size_t)(m0 + i) * ldy + r], t);
            }
        }
    }
}

// AVX-512 VNNI kernel: two Q40 blocks (64 weights) per instruction.  Activations are re-packed once per call into
// contiguous int8 pairs of blocks + a 16-lane vector of their scales, so the inner loop is: one 64-byte load, a masked
// negate (sign transfer), VPDPBUSD, convert, scale, accumulate.  One accumulator chain per row for every batch size, so
// results do not depend on the batch a row travels in (same property as the AVX2 kernel).
__attribute__((target("avx512f,avx512bw,avx512vl,avx512vnni,avx2,fma,f16c")))
void matmul_q40_q80_vnni(const uint8_t* w, int r0, int r1, int nb, const uint8_t* xq, int M, float* y, size_t ldy) {
    const int np = (nb + 1) / 2;
    // [M][np][64] int8 and [M][np][16] f32 (lanes 0-7: scale of the even block, 8-15: of the odd block; padding = 0)
    std::vector<int8_t> x8((size_t)M * np * 64 + 64, 0);
    std::vector<float> xs((size_t)M * np * 16 + 16, 0.f);
    for (int m = 0; m < M; m++)
        for (int b = 0; b < nb; b++) {
            const uint8_t* xb = xq + ((size_t)m * nb + b) * Q80_BYTES;
            uint16_t h;
            memcpy(&h, xb, 2);
            const float d = _cvtsh_ss(h);
            const size_t pi = (size_t)m * np + b / 2;
            memcpy(&x8[pi * 64 + (b & 1) * 32], xb + 2, 32);
            float* sv = &xs[pi * 16 + (b & 1) * 8];
            for (int j = 0; j < 8; j++) sv[j] = d;
        }
    const __m512i low = _mm512_set1_epi8(0x0f), eight = _mm512_set1_epi8(8), zero = _mm512_setzero_si512();
    constexpr int MB = 8;
    for (int r = r0; r < r1; r++) {
        const uint8_t* wr = w + (size_t)r * nb * Q40_BYTES;
        for (int m0 = 0; m0 < M; m0 += MB) {
            const int mc = std::min(MB, M - m0);
            __m512 acc[MB];
            for (int i = 0; i < MB; i++) acc[i] = _mm512_setzero_ps();
            for (int p = 0; p < np; p++) {
                const uint8_t* w0 = wr + (size_t)(2 * p) * Q40_BYTES;
                const bool pair = 2 * p + 1 < nb;
                uint16_t h0, h1 = 0;
                memcpy(&h0, w0, 2);
                const __m128i raw0 = _mm_loadu_si128((const __m128i*)(w0 + 2));
                __m128i raw1 = _mm_setzero_si128();
                if (pair) {
                    memcpy(&h1, w0 + Q40_BYTES, 2);
                    raw1 = _mm_loadu_si128((const __m128i*)(w0 + Q40_BYTES + 2));
                }
                const __m256i v0 = _mm256_set_m128i(_mm_srli_epi16(raw0, 4), raw0);      // low nibbles = elements 0..15
                const __m256i v1 = _mm256_set_m128i(_mm_srli_epi16(raw1, 4), raw1);
                __m512i wv = _mm512_inserti64x4(_mm512_castsi256_si512(v0), v1, 1);
                wv = _mm512_sub_epi8(_mm512_and_si512(wv, low), eight);
                const __m512i aw = _mm512_abs_epi8(wv);
                const __mmask64 neg = _mm512_movepi8_mask(wv);
                const __m512 dwv = _mm512_mask_blend_ps((__mmask16)0xFF00, _mm512_set1_ps(_cvtsh_ss(h0)),
                                                        _mm512_set1_ps(pair ? _cvtsh_ss(h1) : 0.f));
                for (int i = 0; i < mc; i++) {
                    const size_t pi = (size_t)(m0 + i) * np + p;
                    const __m512i xv = _mm512_loadu_si512((const void*)&x8[pi * 64]);
                    const __m512i sx = _mm512_mask_sub_epi8(xv, neg, zero, xv);            // x with the weight's sign
                    const __m512 pf = _mm512_cvtepi32_ps(_mm512_dpbusd_epi32(zero, aw, sx));
                    acc[i] = _mm512_fmadd_ps(pf, _mm512_mul_ps(dwv, _mm512_loadu_ps(&xs[pi * 16])), acc[i]);
                }
            }
            for (int i = 0; i < mc; i++) y[(size_t)(m0 + i) * ldy + r] = _mm512_reduce_add_ps(acc[i]);
        }
    }
}

__attribute__((target("avx2,fma")))
float dot_f32_avx2(const float* a, const float* b, int n) {
    __m256 acc = _mm256_setzero_ps();
    int i = 0;
    for (; i + 8 <= n; i += 8) acc = _mm256_fmadd_ps(_mm256_loadu_ps(a + i), _mm256_loadu_ps(b + i), acc);
    __m128 s = _mm_add_ps(_mm256_extractf128_ps(acc, 1), _mm256_castps256_ps128(acc));
    s = _mm_add_ps(s, _mm_movehl_ps(s, s));
    s = _mm_add_ss(s, _mm_movehdup_ps(s));
    float r = _mm_cvtss_f32(s);
    for (; i < n; i++) r += a[i] * b[i];
    return r;
}
#endif

bool has_vnni() {
#ifdef DL_X86
    static const bool v = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") &&
                          __builtin_cpu_supports("avx512vl") && __builtin_cpu_supports("avx512vnni") &&
                          __builtin_cpu_supports("f16c") && !getenv("DLLAMA_CPU_SCALAR") && !getenv("DLLAMA_CPU_NO_AVX512");
    return v;
#else
    return false;
#endif
}

bool has_avx2() {
#ifdef DL_X86
    static const bool v = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("fma") && __builtin_cpu_supports("f16c")
                          && !getenv("DLLAMA_CPU_SCALAR");
    return v;
#else
    return false;
#endif
}

void quantize_q80(const float* x, uint8_t* out, int n) {
#ifdef DL_X86
    if (has_avx2()) return quantize_q80_avx2(x, out, n);
#endif
    quantize_q80_scalar(x, out, n);
}

void matmul_q40_q80(const uint8_t* w, int r0, int r1, int nb, const uint8_t* xq, int M, float* y, size_t ldy) {
#ifdef DL_X86
    if (has_vnni()) return matmul_q40_q80_vnni(w, r0, r1, nb, xq, M, y, ldy);
    if (has_avx2()) return matmul_q40_q80_avx2(w, r0, r1, nb, xq, M, y, ldy);
#endif
    matmul_q40_q80_scalar(w, r0, r1, nb, xq, M, y, ldy);
}

#ifdef DL_X86
__attribute__((target("avx2,fma")))
void axpy_f32_avx2(float a, const float* x, float* y, int n) {
    const __m256 av = _mm256_set1_ps(a);
    int i = 0;
    for (; i + 8 <= n; i += 8) _mm256_storeu_ps(y + i, _mm256_fmadd_ps(av, _mm256_loadu_ps(x + i), _mm256_loadu_ps(y + i)));
    for (; i < n; i++) y[i] += a * x[i];
}
#endif

#ifdef DL_X86
// exp(x) for x <= 0 (softmax arguments after max subtraction), 8 lanes: range reduction x = n ln2 + r, degree-6
// polynomial on r in [-ln2/2, ln2/2] (Cephes coefficients), relative error ~1e-7; inputs below -87 give 0
// (reference: the custom exp polynomials of its AVX2 / NEON softmax, nn-cpu-ops.cpp:82-102)
__attribute__((target("avx2,fma")))
static inline __m256 exp8_neg(__m256 x) {
    const __m256 lo = _mm256_set1_ps(-87.0f);
    const __m256 under = _mm256_cmp_ps(x, lo, _CMP_LT_OQ);
    x = _mm256_max_ps(x, lo);
    __m256 n = _mm256_round_ps(_mm256_mul_ps(x, _mm256_set1_ps(1.44269504088896341f)), _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC);
    __m256 r = _mm256_fnmadd_ps(n, _mm256_set1_ps(0.693359375f), x);
    r = _mm256_fnmadd_ps(n, _mm256_set1_ps(-2.12194440e-4f), r);
    __m256 p = _mm256_set1_ps(1.9875691500e-4f);
    p = _mm256_fmadd_ps(p, r, _mm256_set1_ps(1.3981999507e-3f));
    p = _mm256_fmadd_ps(p, r, _mm256_set1_ps(8.3334519073e-3f));
    p = _mm256_fmadd_ps(p, r, _mm256_set1_ps(4.1665795894e-2f));
    p = _mm256_fmadd_ps(p, r, _mm256_set1_ps(1.6666665459e-1f));
    p = _mm256_fmadd_ps(p, r, _mm256_set1_ps(5.0000001201e-1f));
    p = _mm256_fmadd_ps(p, _mm256_mul_ps(r, r), _mm256_add_ps(r, _mm256_set1_ps(1.0f)));
    const __m256i e = _mm256_slli_epi32(_mm256_add_epi32(_mm256_cvtps_epi32(n), _mm256_set1_epi32(127)), 23);
    return _mm256_andnot_ps(under, _mm256_mul_ps(p, _mm256_castsi256_ps(e)));
}

// s[0..n) <- softmax(s[0..n)) in place; elements are processed 8 at a time, the tail with the same polynomial through
// a padded register, sums in lane order -- deterministic for a given n
__attribute__((target("avx2,fma")))
void softmax_inplace_avx2(float* s, int n) {
    float mx = -INFINITY;
    for (int i = 0; i < n; i++) mx = std::max(mx, s[i]);
    const __m256 mv = _mm256_set1_ps(mx);
    __m256 acc = _mm256_setzero_ps();
    int i = 0;
    for (; i + 8 <= n; i += 8) {
        const __m256 e = exp8_neg(_mm256_sub_ps(_mm256_loadu_ps(s + i), mv));
        _mm256_storeu_ps(s + i, e);
        acc = _mm256_add_ps(acc, e);
    }
    float sum;
    DL_HSUM8(sum, acc);
    if (i < n) {
        float tmp[8];
        for (int k = 0; k < 8; k++) tmp[k] = i + k < n ? s[i + k] : -INFINITY;
        const __m256 e = exp8_neg(_mm256_sub_ps(_mm256_loadu_ps(tmp), mv));
        _mm256_storeu_ps(tmp, e);
        for (int k = 0; i + k < n; k++) { s[i + k] = tmp[k]; sum += tmp[k]; }
    }
    const float inv = 1.0f / sum;
    const __m256 iv = _mm256_set1_ps(inv);
    for (i = 0; i + 8 <= n; i += 8) _mm256_storeu_ps(s + i, _mm256_mul_ps(_mm256_loadu_ps(s + i), iv));
    for (; i < n; i++) s[i] *= inv;
}
#endif

// softmax with max subtraction (reference: softmax_F32, nn-cpu-ops.cpp)
void softmax_inplace(float* s, int n) {
#ifdef DL_X86
    if (has_avx2()) return softmax_inplace_avx2(s, n);
#endif
    float mx = -INFINITY;
    for (int i = 0; i < n; i++) mx = std::max(mx, s[i]);
    float sum = 0.f;
    for (int i = 0; i < n; i++) { s[i] = expf(s[i] - mx); sum += s[i]; }
    const float inv = 1.0f / sum;
    for (int i = 0; i < n; i++) s[i] *= inv;
}

#ifdef DL_X86
// out[g] = scale * dot(q + g*n, k) for g < rep: the K row is loaded once for four query heads at a time and the four
// accumulators are reduced together (two hadd levels) instead of one horizontal sum per dot
__attribute__((target("avx2,fma")))
void dots_avx2(const float* q, const float* k, int n, int rep, float scale, float* out, size_t out_stride) {
    int g = 0;
    for (; g + 4 <= rep; g += 4) {
        const float* q0 = q + (size_t)g * n;
        __m256 a0 = _mm256_setzero_ps(), a1 = a0, a2 = a0, a3 = a0;
        for (int i = 0; i + 8 <= n; i += 8) {
            const __m256 kv = _mm256_loadu_ps(k + i);
            a0 = _mm256_fmadd_ps(_mm256_loadu_ps(q0 + i), kv, a0);
            a1 = _mm256_fmadd_ps(_mm256_loadu_ps(q0 + n + i), kv, a1);
            a2 = _mm256_fmadd_ps(_mm256_loadu_ps(q0 + 2 * n + i), kv, a2);
            a3 = _mm256_fmadd_ps(_mm256_loadu_ps(q0 + 3 * n + i), kv, a3);
        }
        const __m256 h = _mm256_hadd_ps(_mm256_hadd_ps(a0, a1), _mm256_hadd_ps(a2, a3));     // per 128-bit lane: sums of a0..a3
        __m128 r = _mm_add_ps(_mm256_castps256_ps128(h), _mm256_extractf128_ps(h, 1));
        r = _mm_mul_ps(r, _mm_set1_ps(scale));
        float t[4];
        _mm_storeu_ps(t, r);
        for (int j = 0; j < 4; j++) out[(size_t)(g + j) * out_stride] = t[j];
    }
    for (; g < rep; g++) out[(size_t)g * out_stride] = dot_f32_avx2(q + (size_t)g * n, k, n) * scale;
}
#endif

void axpy_f32(float a, const float* x, float* y, int n);

#ifdef DL_X86
// o[0..n) = sum_p w[p] * v[p * v_stride + 0..n): the output stays in registers (64 floats at a time) over the whole
// position loop instead of being loaded and stored once per position
__attribute__((target("avx2,fma")))
void weighted_sum_avx2(const float* w, const float* v, size_t v_stride, int count, float* o, int n) {
    for (int c0 = 0; c0 < n; c0 += 64) {
        __m256 a0 = _mm256_setzero_ps(), a1 = a0, a2 = a0, a3 = a0, a4 = a0, a5 = a0, a6 = a0, a7 = a0;
        for (int p = 0; p < count; p++) {
            const __m256 wp = _mm256_set1_ps(w[p]);
            const float* vr = v + (size_t)p * v_stride + c0;
            a0 = _mm256_fmadd_ps(wp, _mm256_loadu_ps(vr), a0);
            a1 = _mm256_fmadd_ps(wp, _mm256_loadu_ps(vr + 8), a1);
            a2 = _mm256_fmadd_ps(wp, _mm256_loadu_ps(vr + 16), a2);
            a3 = _mm256_fmadd_ps(wp, _mm256_loadu_ps(vr + 24), a3);
            a4 = _mm256_fmadd_ps(wp, _mm256_loadu_ps(vr + 32), a4);
            a5 = _mm256_fmadd_ps(wp, _mm256_loadu_ps(vr + 40), a5);
            a6 = _mm256_fmadd_ps(wp, _mm256_loadu_ps(vr + 48), a6);
            a7 = _mm256_fmadd_ps(wp, _mm256_loadu_ps(vr + 56), a7);
        }
        float* oc = o + c0;
        _mm256_storeu_ps(oc, a0); _mm256_storeu_ps(oc + 8, a1); _mm256_storeu_ps(oc + 16, a2); _mm256_storeu_ps(oc + 24, a3);
        _mm256_storeu_ps(oc + 32, a4); _mm256_storeu_ps(oc + 40, a5); _mm256_storeu_ps(oc + 48, a6); _mm256_storeu_ps(oc + 56, a7);
    }
}
#endif

// o = sum_p w[p] * v[p]   (the softmax-weighted sum of the cached V rows of one head)
void weighted_sum_f32(const float* w, const float* v, size_t v_stride, int count, float* o, int n) {
#ifdef DL_X86
    if (has_avx2() && n % 64 == 0) return weighted_sum_avx2(w, v, v_stride, count, o, n);
#endif
    for (int i = 0; i < n; i++) o[i] = 0.f;
    for (int p = 0; p < count; p++) axpy_f32(w[p], v + (size_t)p * v_stride, o, n);
}

void dots_f32(const float* q, const float* k, int n, int rep, float scale, float* out, size_t out_stride) {
#ifdef DL_X86
    if (has_avx2() && n % 8 == 0) return dots_avx2(q, k, n, rep, scale, out, out_stride);
#endif
    for (int g = 0; g < rep; g++) {
        float r = 0.f;
        for (int i = 0; i < n; i++) r += q[(size_t)g * n + i] * k[i];
        out[(size_t)g * out_stride] = r * scale;
    }
}

// y += a * x
void axpy_f32(float a, const float* x, float* y, int n) {
#ifdef DL_X86
    if (has_avx2()) return axpy_f32_avx2(a, x, y, n);
#endif
    for (int i = 0; i < n; i++) y[i] += a * x[i];
}

float dot_f32(const float* a, const float* b, int n) {
#ifdef DL_X86
    if (has_avx2()) return dot_f32_avx2(a, b, n);
#endif
    float r = 0.f;
    for (int i = 0; i < n; i++) r += a[i] * b[i];
    return r;
}

// ---- element-wise op kernels (one row each; the Node methods below split rows / elements over the pool) ------------
// 1 / sqrt(mean(x^2) + eps)   (reference: invRms_F32, nn-cpu-ops.cpp:105-134)
float inv_rms(const float* x, int n, float eps) { return 1.0f / sqrtf(dot_f32(x, x, n) / (float)n + eps); }

// y = w * (inv * x)   (reference: rmsNorm_F32, nn-cpu-ops.cpp:136-180)
void rms_norm_row(const float* x, const float* w, float inv, float* y, int n) {
    for (int i = 0; i < n; i++) y[i] = w[i] * (inv * x[i]);
}

// SiLU (act == 1) or tanh-GELU (act == 0) in place   (reference: silu_F32 / gelu_F32, nn-cpu-ops.cpp:672-720)
inline float activate(float g, int act) {
    if (act == 1) return g / (1.0f + expf(-g));
    return 0.5f * g * (1.0f + tanhf(0.797884560802865f * (g + 0.044715f * g * g * g)));
}

// rotate adjacent pairs of every head of one row at position `pos`   (reference: ropeLlamaForward_F32_F32,
// nn-cpu-ops.cpp:1090-1120; freq[j] is the rotation frequency of pair j of a head)
void rope_row(float* r, int width, int head_size, int pos, const float* freq) {
    for (int i = 0; i < width; i += 2) {
        const float ang = (float)pos * freq[(i % head_size) / 2];
        const float cs = cosf(ang), sn = sinf(ang);
        const float a = r[i], b = r[i + 1];
        r[i] = a * cs - b * sn;
        r[i + 1] = a * sn + b * cs;
    }
}

// rotation frequency of pair j of a head, with the Llama-3.1 wavelength-dependent scaling when factor != 1
// (reference: fullfillRopeLlama3Cache + scaleFrequencyLlama3, nn-core.cpp:307-340)
float rope_frequency(int j, int head_size, float theta, int rope_type, float factor, float low, float high, int orig_len) {
    float freq = 1.0f / powf(theta, (float)(2 * j) / (float)head_size);
    if (rope_type == 2 && factor != 1.0f) {
        const float wave = 2.0f * (float)M_PI / freq;
        const float high_wl = (float)orig_len / high, low_wl = (float)orig_len / low;
        if (wave < high_wl) {
        } else if (wave > low_wl) freq = freq / factor;
        else {
            const float smooth = ((float)orig_len / wave - low) / (high - low);
            freq = (1 - smooth) * freq / factor + smooth * freq;
        }
    }
    return freq;
}

// ---- persistent thread pool (reference: NnExecutor spawns nThreads-1 pthreads per forward and spins on two
// atomics, nn-executor.cpp:134-187).  Workers spin for a few tens of microseconds between parallel regions (a forward
// pass is hundreds of back-to-back regions) and fall asleep on a condition variable when the engine goes idle. -------
inline void cpu_relax() {
#ifdef DL_X86
    _mm_pause();
#endif
}

class Pool {
public:
    explicit Pool(int n) : n_(std::max(1, n)) {
        for (int t = 1; t < n_; t++) workers_.emplace_back([this, t] { loop(t); });
    }
    ~Pool() {
        quit_.store(true);
        gen_.fetch_add(1);
        {
            std::lock_guard<std::mutex> g(mu_);
        }
        cv_.notify_all();
        for (auto& w : workers_) w.join();
    }
    int size() const { return n_; }
    // fn(thread, begin, end) over [0, total) split evenly (reference: SPLIT_THREADS, nn-quants.hpp:82-86)
    void run(int total, const std::function<void(int, int, int)>& fn) {
        if (n_ == 1 || total < 2) { fn(0, 0, total); return; }
        fn_ = &fn;
        total_ = total;
        pending_.store(n_ - 1);
        gen_.fetch_add(1);                      // seq_cst: published before the sleepers check below
        if (sleepers_.load() > 0) {
            { std::lock_guard<std::mutex> g(mu_); }
            cv_.notify_all();
        }
        work(0);
        int spins = 0;
        while (pending_.load(std::memory_order_acquire) != 0) {
            if (++spins < 4096) cpu_relax(); else std::this_thread::yield();
        }
    }

private:
    void work(int t) {
        const int per = total_ / n_, rem = total_ % n_;
        const int b = t * per + std::min(t, rem), e = b + per + (t < rem ? 1 : 0);
        if (e > b) (*fn_)(t, b, e);
    }
    void loop(int t) {
        static const int kSpin = getenv("DLLAMA_CPU_SPIN") ? atoi(getenv("DLLAMA_CPU_SPIN")) : 200000;
        unsigned long seen = 0;
        while (true) {
            int spins = 0;
            while (gen_.load(std::memory_order_acquire) == seen) {
                if (++spins < kSpin) { cpu_relax(); continue; }
                std::unique_lock<std::mutex> g(mu_);
                sleepers_.fetch_add(1);
                cv_.wait(g, [&] { return gen_.load() != seen; });
                sleepers_.fetch_sub(1);
            }
            seen = gen_.load(std::memory_order_acquire);
            if (quit_.load()) return;
            work(t);
            pending_.fetch_sub(1, std::memory_order_release);
        }
    }
    int n_;
    std::vector<std::thread> workers_;
    std::mutex mu_;
    std::condition_variable cv_;
    const std::function<void(int, int, int)>* fn_ = nullptr;
    int total_ = 0;
    std::atomic<int> pending_{0}, sleepers_{0};
    std::atomic<unsigned long> gen_{0};
    std::atomic<bool> quit_{false};
};

// ---- model description ------------------------------------------------------------------------------
struct Header {                                 // fixed-width: also the wire format of the config broadcast
    int32_t dim, hidden_dim, n_layers, n_heads, n_kv_heads, vocab, seq_len, hidden_act, rope_type, weight_type;
    float rope_theta, rope_factor, rope_low, rope_high, norm_eps;
    int32_t rope_orig_len;
    int64_t header_size, file_size;
    int head_size() const { return dim / n_heads; }
    int kv_dim() const { return dim * n_kv_heads / n_heads; }
};

struct MappedFile {                             // (reference: src/mmap.hpp:35-81)
    int fd = -1;
    const uint8_t* data = nullptr;
    size_t size = 0;
    bool open(const char* path) {
        fd = ::open(path, O_RDONLY);
        if (fd < 0) return false;
        struct stat st;
        if (fstat(fd, &st) != 0) return false;
        size = (size_t)st.st_size;
        void* p = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (p == MAP_FAILED) return false;
        data = (const uint8_t*)p;
        return true;
    }
    ~MappedFile() {
        if (data) munmap((void*)data, size);
        if (fd >= 0) close(fd);
    }
};

// (reference: loadLlmHeader, src/llm.cpp:26-98; keys src/llm.hpp:8-28)
void load_header(const MappedFile& f, int max_seq_len, Header& h) {
    if (f.size < 8) throw std::runtime_error("Cannot read magic value");
    int32_t magic, hsize;
    memcpy(&magic, f.data, 4);
    memcpy(&hsize, f.data + 4, 4);
    if (magic == 0xABCD00 || magic == 0xABCD01) throw std::runtime_error("Old model format is not supported");
    if (magic != 0xA00ABCD) throw std::runtime_error("Unsupported magic number");
    if (hsize < 8 || (size_t)hsize > f.size) throw std::runtime_error("Cannot read header values");
    memset(&h, 0, sizeof h);
    h.hidden_act = 1;
    h.rope_theta = 10000.f;
    h.rope_factor = 1.f;
    h.norm_eps = 1e-5f;
    h.weight_type = -1;
    int arch = 0xABCD00;
    const int n = (hsize - 8) / 4;
    for (int i = 0; i + 1 < n; i += 2) {
        int32_t k, v;
        memcpy(&k, f.data + 8 + 4 * i, 4);
        memcpy(&v, f.data + 12 + 4 * i, 4);
        switch (k) {
            case 0: break;
            case 1: arch = v; break;
            case 2: h.dim = v; break;
            case 3: h.hidden_dim = v; break;
            case 4: h.n_layers = v; break;
            case 5: h.n_heads = v; break;
            case 6: h.n_kv_heads = v; break;
            case 7: case 8: break;
            case 9: h.vocab = v; break;
            case 10: h.seq_len = v; break;
            case 11: h.hidden_act = v; break;
            case 12: h.rope_theta = (float)v; break;
            case 13: h.weight_type = v; break;
            case 14: h.rope_factor = (float)v; break;
            case 15: h.rope_low = (float)v; break;
            case 16: h.rope_high = (float)v; break;
            case 17: h.rope_orig_len = v; break;
            case 18: h.rope_type = v; break;
            default: throw std::runtime_error("Unsupported header key");
        }
    }
    if (h.weight_type < 0) throw std::runtime_error("Model does not specify weight type");
    if (arch != 0xABCD00) throw std::runtime_error("Unsupported architecture");
    if (max_seq_len > 0 && h.seq_len > max_seq_len) h.seq_len = max_seq_len;
    h.header_size = hsize;
    h.file_size = (int64_t)f.size;
}

size_t row_bytes(int ft, int n) {
    switch (ft) {
        case F_32: return (size_t)n * 4;
        case F_16: return (size_t)n * 2;
        case F_Q40: return (size_t)(n / QK) * Q40_BYTES;
        case F_Q80: return (size_t)(n / QK) * Q80_BYTES;
    }
    throw std::runtime_error("Unsupported float type");
}

// One matrix slice owned by a node: Q40 blocks as in the file, or f32 (f16 files are widened at load).
struct Mat {
    int d = 0, n = 0;                  // local output rows, local input columns
    bool q40 = false;
    std::vector<uint8_t> q;            // [d][n/32] Q40 blocks
    std::vector<float> f;              // [d][n] f32
};

// rows [r0, r0+d) of a [D][n] matrix (output split; reference: splitRowMatmulWeight, nn-core.cpp:270-286)
void slice_rows(const uint8_t* src, int ft, int n, int r0, int d, Mat& out) {
    out.d = d;
    out.n = n;
    const size_t rb = row_bytes(ft, n);
    if (ft == F_Q40) {
        out.q40 = true;
        out.q.assign(src + (size_t)r0 * rb, src + (size_t)(r0 + d) * rb);
    } else {
        out.f.resize((size_t)d * n);
        for (int r = 0; r < d; r++) {
            const uint8_t* p = src + (size_t)(r0 + r) * rb;
            if (ft == F_32) memcpy(&out.f[(size_t)r * n], p, (size_t)n * 4);
            else for (int c = 0; c < n; c++) { uint16_t v; memcpy(&v, p + 2 * c, 2); out.f[(size_t)r * n + c] = f16_to_f32(v); }
        }
    }
}

// columns [c0, c0+n) of a [d][N] matrix (input split, whole quant blocks; reference: splitColMatmulWeight, :288-303)
void slice_cols(const uint8_t* src, int ft, int N, int d, int c0, int n, Mat& out) {
    out.d = d;
    out.n = n;
    const size_t rb = row_bytes(ft, N);
    if (ft == F_Q40) {
        if (c0 % QK || n % QK) throw std::runtime_error("column slice is not a multiple of the quant block");
        out.q40 = true;
        const size_t sb = (size_t)(n / QK) * Q40_BYTES;
        out.q.resize((size_t)d * sb);
        for (int r = 0; r < d; r++) memcpy(&out.q[(size_t)r * sb], src + (size_t)r * rb + (size_t)(c0 / QK) * Q40_BYTES, sb);
    } else {
        out.f.resize((size_t)d * n);
        for (int r = 0; r < d; r++) {
            const uint8_t* p = src + (size_t)r * rb;
            if (ft == F_32) memcpy(&out.f[(size_t)r * n], p + (size_t)c0 * 4, (size_t)n * 4);
            else for (int c = 0; c < n; c++) { uint16_t v; memcpy(&v, p + 2 * (c0 + c), 2); out.f[(size_t)r * n + c] = f16_to_f32(v); }
        }
    }
}

struct LayerW { Mat wq, wk, wv, wo, w1, w2, w3; std::vector<float> rms_att, rms_ffn; };

struct NodeConfig {                  // what the root tells a worker (reference: NnNetConfig + NnNodeConfig broadcast)
    Header h;
    int32_t n_nodes, node, max_rows, n_slots, sync_q80, n_threads;
};

// ---- sockets ------------------------------------------------------------------------------------------
struct NetStats { std::atomic<long long> sent{0}, recv{0}; };

void set_nodelay(int fd) {
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
}

void send_all(int fd, const void* p, size_t n, NetStats* st) {
    const uint8_t* b = (const uint8_t*)p;
    size_t off = 0;
    while (off < n) {
        ssize_t w = send(fd, b + off, n - off, MSG_NOSIGNAL);
        if (w < 0) {
            if (errno == EINTR) continue;
            if (errno == EAGAIN || errno == EWOULDBLOCK) { pollfd pf{fd, POLLOUT, 0}; poll(&pf, 1, 1000); continue; }
            throw std::runtime_error(std::string("socket write failed: ") + strerror(errno));
        }
        off += (size_t)w;
    }
    if (st) st->sent += (long long)n;
}

void recv_all(int fd, void* p, size_t n, NetStats* st) {
    uint8_t* b = (uint8_t*)p;
    size_t off = 0;
    while (off < n) {
        ssize_t r = recv(fd, b + off, n - off, 0);
        if (r == 0) throw std::runtime_error("socket closed by peer");
        if (r < 0) {
            if (errno == EINTR) continue;
            if (errno == EAGAIN || errno == EWOULDBLOCK) { pollfd pf{fd, POLLIN, 0}; poll(&pf, 1, 1000); continue; }
            throw std::runtime_error(std::string("socket read failed: ") + strerror(errno));
        }
        off += (size_t)r;
    }
    if (st) st->recv += (long long)n;
}

int connect_to(const std::string& host, int port, int timeout_s) {
    addrinfo hints{}, *res = nullptr;
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    if (getaddrinfo(host.c_str(), std::to_string(port).c_str(), &hints, &res) != 0 || !res)
        throw std::runtime_error("cannot resolve " + host);
    const auto t0 = std::chrono::steady_clock::now();
    while (true) {
        int fd = socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
        if (fd >= 0 && connect(fd, res->ai_addr, res->ai_addrlen) == 0) {
            freeaddrinfo(res);
            set_nodelay(fd);
            return fd;
        }
        if (fd >= 0) close(fd);
        if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(timeout_s)) {
            freeaddrinfo(res);
            throw std::runtime_error("cannot connect to " + host + ":" + std::to_string(port));
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
    }
}

// Full mesh between the nodes (reference: NnNetwork::connect / serve, nn-network.cpp:264-348): peer[i] is the
// socket to node i (-1 for self).
struct Mesh {
    int n_nodes = 1, node = 0;
    std::vector<int> peer;
    NetStats stats;
    double sync_s = 0.0;
    ~Mesh() { for (int fd : peer) if (fd >= 0) close(fd); }

    // Every node sends `bytes` from send_buf to each listed peer and receives `bytes` from each into recv_ptr[peer];
    // all transfers progress together under poll(), so nobody blocks on a full socket buffer.
    void exchange(const std::vector<int>& send_to, const uint8_t* send_buf, size_t send_bytes,
                  const std::vector<int>& recv_from, const std::vector<uint8_t*>& recv_ptr, size_t recv_bytes) {
        const auto t0 = std::chrono::steady_clock::now();
        struct Xfer { int fd; bool out; size_t off, total; const uint8_t* src; uint8_t* dst; };
        std::vector<Xfer> x;
        for (int p : send_to) x.push_back({peer[p], true, 0, send_bytes, send_buf, nullptr});
        for (size_t i = 0; i < recv_from.size(); i++) x.push_back({peer[recv_from[i]], false, 0, recv_bytes, nullptr, recv_ptr[i]});
        std::vector<pollfd> pf(x.size());
        size_t left = 0;
        int idle_polls = 0;
        static const int kSpinPolls = getenv("DLLAMA_CPU_SPIN_POLLS") ? atoi(getenv("DLLAMA_CPU_SPIN_POLLS")) : 20000;
        for (auto& t : x) if (t.total > 0) left++;
        while (left > 0) {
            for (size_t i = 0; i < x.size(); i++) {
                pf[i].fd = x[i].off < x[i].total ? x[i].fd : -1;
                pf[i].events = x[i].out ? POLLOUT : POLLIN;
                pf[i].revents = 0;
            }
            // peers reach a synchronisation point within microseconds of each other: look a few times before sleeping
            int rc = poll(pf.data(), pf.size(), idle_polls < kSpinPolls ? 0 : 30000);
            if (rc < 0 && errno == EINTR) continue;
            if (rc == 0 && idle_polls < kSpinPolls) { idle_polls++; cpu_relax(); continue; }
            if (rc <= 0) throw std::runtime_error("node synchronisation timed out");
            idle_polls = 0;
            for (size_t i = 0; i < x.size(); i++) {
                Xfer& t = x[i];
                if (pf[i].fd < 0 || !pf[i].revents) continue;
                if (pf[i].revents & (POLLERR | POLLNVAL)) throw std::runtime_error("socket error during synchronisation");
                ssize_t k = t.out ? send(t.fd, t.src + t.off, t.total - t.off, MSG_NOSIGNAL | MSG_DONTWAIT)
                                  : recv(t.fd, t.dst + t.off, t.total - t.off, MSG_DONTWAIT);
                if (k == 0 && !t.out) throw std::runtime_error("socket closed by peer");
                if (k < 0) {
                    if (errno == EAGAIN || errno == EWOULDBLOCK || errno == EINTR) continue;
                    throw std::runtime_error(std::string("socket transfer failed: ") + strerror(errno));
                }
                t.off += (size_t)k;
                (t.out ? stats.sent : stats.recv) += k;
                if (t.off == t.total) left--;
            }
        }
        sync_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
};

// ---- one node's engine ------------------------------------------------------------------------------------
struct Node {
    NodeConfig c{};
    Mesh mesh;
    std::unique_ptr<Pool> pool;
    // weights of this node
    std::vector<float> embedding;               // root only
    std::vector<LayerW> layers;
    std::vector<float> rms_final;
    Mat wcls;
    std::vector<float> rope_freq;               // [head_size/2]
    // per-slot KV caches, allocated on first use: [slot][layer] -> [seq_len][kv_dim/N]
    std::vector<std::vector<std::vector<float>>> kc, vc;
    // activations
    std::vector<float> x, y, q, k, v, att_out, part, hb, hl, lg;
    std::vector<uint8_t> yq, zq;                // Q80 casts; zq: [n_nodes][M][slice bytes]
    std::vector<std::vector<float>> scores;     // per thread: [rep][seq_len], grown on demand
    std::vector<int> pos, slot;
    // DLLAMA_CPU_PROFILE=1: seconds per op family, printed when the node is destroyed (reference: the executor's
    // per-step-type timers, nn-executor.cpp:150-154)
    bool profile = getenv("DLLAMA_CPU_PROFILE") != nullptr;
    // DLLAMA_CPU_DEBUG_OPS=1: first values of every op's output for row 0 (reference: DEBUG_OP_INPUT_OUTPUT,
    // nn-cpu-ops.cpp:14-27) -- lines can be diffed between node counts, kernels (DLLAMA_CPU_SCALAR) or engines
    bool debug_ops = getenv("DLLAMA_CPU_DEBUG_OPS") != nullptr;
    void dbg(const char* op, int layer, const float* v, int n) const {
        if (!debug_ops) return;
        fprintf(stderr, "[op] node %d layer %2d %-10s", c.node, layer, op);
        for (int i = 0; i < std::min(n, 4); i++) fprintf(stderr, " %+.6f", v[i]);
        fprintf(stderr, "\n");
    }
    double t_matmul = 0, t_att = 0, t_sync = 0, t_other = 0;
    long long n_forward = 0;
    ~Node() {
        if (profile && n_forward > 0)
            fprintf(stderr, "[cpu-profile] node %d: %lld forwards, per forward: matmul %.2f ms, attention %.2f ms, "
                    "sync+merge %.2f ms, other %.2f ms\n", c.node, n_forward, 1e3 * t_matmul / n_forward,
                    1e3 * t_att / n_forward, 1e3 * t_sync / n_forward, 1e3 * t_other / n_forward);
    }

    int N() const { return c.n_nodes; }
    bool q80() const { return c.sync_q80 != 0; }
    size_t zslice_bytes() const { return q80() ? row_bytes(F_Q80, c.h.dim) : (size_t)c.h.dim * 4; }

    void init_buffers() {
        const Header& h = c.h;
        const int R = c.max_rows, n = N();
        pool.reset(new Pool(c.n_threads));
        x.resize((size_t)R * h.dim);
        y.resize((size_t)R * h.dim);
        part.resize((size_t)R * h.dim);
        q.resize((size_t)R * h.dim / n);
        k.resize((size_t)R * h.kv_dim() / n);
        v.resize((size_t)R * h.kv_dim() / n);
        att_out.resize((size_t)R * h.dim / n);
        hb.resize((size_t)R * h.hidden_dim / n);
        hl.resize((size_t)R * h.hidden_dim / n);
        lg.resize((size_t)R * h.vocab / n);
        yq.resize((size_t)R * row_bytes(F_Q80, std::max(h.dim, h.hidden_dim)));
        zq.resize((size_t)n * R * zslice_bytes());
        scores.assign(pool->size(), std::vector<float>(h.seq_len));
        kc.assign(c.n_slots, {});
        vc.assign(c.n_slots, {});
        pos.resize(R);
        slot.resize(R);
        // rotation frequency of pair j of a head (reference: fullfillRopeLlama3Cache, nn-core.cpp:307-340)
        const int hs = h.head_size();
        rope_freq.resize(hs / 2);
        for (int j = 0; j < hs / 2; j++)
            rope_freq[j] = rope_frequency(j, hs, h.rope_theta, h.rope_type, h.rope_factor, h.rope_low, h.rope_high, h.rope_orig_len);
    }

    void ensure_slot(int s) {
        if (s < 0 || s >= c.n_slots) throw std::runtime_error("KV slot out of range");
        if (!kc[s].empty()) return;
        const size_t n = (size_t)c.h.seq_len * c.h.kv_dim() / N();
        kc[s].assign(c.h.n_layers, std::vector<float>());
        vc[s].assign(c.h.n_layers, std::vector<float>());
        for (int l = 0; l < c.h.n_layers; l++) { kc[s][l].assign(n, 0.f); vc[s][l].assign(n, 0.f); }
    }

    // ---- ops ----
    // y = w * (x / rms(x))   (reference: OP_INV_RMS + OP_RMS_NORM, nn-cpu-ops.cpp:105-134,944-963)
    void rmsnorm(const float* in, const float* w, float* out, int M, int dim) {
        pool->run(M, [&](int, int b, int e) {
            for (int m = b; m < e; m++) {
                const float* r = in + (size_t)m * dim;
                rms_norm_row(r, w, inv_rms(r, dim, c.h.norm_eps), out + (size_t)m * dim, dim);
            }
        });
    }

    // out[m][r] = W[r] . in[m]; the input is cast to Q80 first when the weights are Q40
    void matmul(const Mat& w, const float* in, int ld_in, int M, float* out, int ldo) {
        if (w.q40) {
            const size_t xb = row_bytes(F_Q80, w.n);
            pool->run(M, [&](int, int b, int e) {
                for (int m = b; m < e; m++) quantize_q80(in + (size_t)m * ld_in, yq.data() + (size_t)m * xb, w.n);
            });
            pool->run(w.d, [&](int, int b, int e) { matmul_q40_q80(w.q.data(), b, e, w.n / QK, yq.data(), M, out, (size_t)ldo); });
        } else {
            pool->run(w.d, [&](int, int b, int e) {
                for (int r = b; r < e; r++)
                    for (int m = 0; m < M; m++)
                        out[(size_t)m * ldo + r] = dot_f32(&w.f[(size_t)r * w.n], in + (size_t)m * ld_in, w.n);
            });
        }
    }

    // rotate adjacent pairs of every head (reference: ropeLlamaForward_F32_F32, nn-cpu-ops.cpp:1090-1120)
    void rope(float* t, int M, int width) {
        const int hs = c.h.head_size();
        pool->run(M, [&](int, int b, int e) {
            for (int m = b; m < e; m++) rope_row(t + (size_t)m * width, width, hs, pos[m], rope_freq.data());
        });
    }

    // causal attention of every (row, local head) over that row's own sequence (reference: multiheadAtt_F32,
    // nn-cpu-ops.cpp:749-784; here per-row position and KV slot)
    void attention(int l, int M) {
        const int hs = c.h.head_size(), nh = c.h.n_heads / N(), nkv = c.h.n_kv_heads / N(), rep = nh / nkv;
        const int kvw = nkv * hs, qw = nh * hs, S = c.h.seq_len;
        const float scale = 1.0f / sqrtf((float)hs);
        // one work item = (row, KV head): the `rep` query heads that share a KV head are processed together, so every
        // cached K and V row is read once per group instead of once per query head (grouped-query attention).
        // (Also sharing the K/V rows between the rows of a prompt chunk was measured and is slower: the score matrix of
        // a whole chunk no longer fits the L1/L2 working set, and the kernel is bound by dot / exp work, not by K/V traffic.)
        pool->run(M * nkv, [&](int t, int b, int e) {
            if (scores[t].size() < (size_t)rep * S) scores[t].resize((size_t)rep * S);
            float* sc = scores[t].data();
            for (int it = b; it < e; it++) {
                const int m = it / nkv, kvh = it % nkv, P = pos[m];
                const float* kk = kc[slot[m]][l].data() + (size_t)kvh * hs;
                const float* vv = vc[slot[m]][l].data() + (size_t)kvh * hs;
                const float* qq = q.data() + (size_t)m * qw + (size_t)kvh * rep * hs;
                for (int p = 0; p <= P; p++) {
                    const float* kr = kk + (size_t)p * kvw;
                    dots_f32(qq, kr, hs, rep, scale, sc + p, (size_t)S);
                }
                for (int g = 0; g < rep; g++) softmax_inplace(sc + (size_t)g * S, P + 1);
                float* o = att_out.data() + (size_t)m * qw + (size_t)kvh * rep * hs;
                for (int g = 0; g < rep; g++) weighted_sum_f32(sc + (size_t)g * S, vv, (size_t)kvw, P + 1, o + (size_t)g * hs, hs);
            }
        });
    }

    // partial sum -> this node's slice of ZQ (cast to the sync type), all-gather, x += sum of the slices
    // (reference: OP_CAST y->ZQ + SYNC_NODE_SLICES + OP_MERGE_ADD, llm.cpp:308-322,381-387)
    void sync_and_merge(int M) {
        const int dim = c.h.dim, n = N();
        const size_t sb = zslice_bytes(), blk = (size_t)M * sb;
        uint8_t* mine = zq.data() + (size_t)c.node * blk;
        pool->run(M, [&](int, int b, int e) {
            for (int m = b; m < e; m++) {
                if (q80()) quantize_q80(part.data() + (size_t)m * dim, mine + (size_t)m * sb, dim);
                else memcpy(mine + (size_t)m * sb, part.data() + (size_t)m * dim, sb);
            }
        });
        if (n > 1) {
            std::vector<int> others;
            std::vector<uint8_t*> dst;
            for (int p = 0; p < n; p++) if (p != c.node) { others.push_back(p); dst.push_back(zq.data() + (size_t)p * blk); }
            mesh.exchange(others, mine, blk, others, dst, blk);
        }
        pool->run(M, [&](int t, int b, int e) {
            std::vector<float> tmp(dim), acc(dim);
            for (int m = b; m < e; m++) {
                for (int p = 0; p < n; p++) {
                    const uint8_t* s = zq.data() + (size_t)p * blk + (size_t)m * sb;
                    const float* f;
                    if (q80()) { dequantize_q80(s, tmp.data(), dim); f = tmp.data(); } else f = (const float*)s;
                    if (p == 0) memcpy(acc.data(), f, (size_t)dim * 4);
                    else for (int i = 0; i < dim; i++) acc[i] += f[i];
                }
                float* xr = x.data() + (size_t)m * dim;
                for (int i = 0; i < dim; i++) xr[i] += acc[i];
            }
        });
    }

    // One forward pass of this node for M rows whose residual stream x is already filled.  Leaves this node's
    // logits slice for rows [lf, M) in lg ([M-lf][vocab/N]).
    void forward_node(int M, int lf) {
        const Header& h = c.h;
        const int n = N(), dim = h.dim, qw = dim / n, kvw = h.kv_dim() / n, hw = h.hidden_dim / n;
        for (int m = 0; m < M; m++) {
            if (pos[m] < 0 || pos[m] >= h.seq_len) throw std::runtime_error("position outside the context window");
            ensure_slot(slot[m]);
        }
        using clk = std::chrono::steady_clock;
        auto t_last = clk::now();
        auto lap = [&](double& acc) {
            if (!profile) return;
            const auto t = clk::now();
            acc += std::chrono::duration<double>(t - t_last).count();
            t_last = t;
        };
        n_forward++;
        for (int l = 0; l < h.n_layers; l++) {
            LayerW& L = layers[l];
            rmsnorm(x.data(), L.rms_att.data(), y.data(), M, dim);
            dbg("rms_att", l, y.data(), dim);
            lap(t_other);
            matmul(L.wq, y.data(), dim, M, q.data(), qw);
            matmul(L.wk, y.data(), dim, M, k.data(), kvw);
            matmul(L.wv, y.data(), dim, M, v.data(), kvw);
            dbg("q", l, q.data(), qw);
            dbg("k", l, k.data(), kvw);
            dbg("v", l, v.data(), kvw);
            lap(t_matmul);
            rope(q.data(), M, qw);
            rope(k.data(), M, kvw);
            dbg("q_rope", l, q.data(), qw);
            for (int m = 0; m < M; m++) {           // (reference: OP_SHIFT, nn-cpu-ops.cpp:1253-1275)
                memcpy(kc[slot[m]][l].data() + (size_t)pos[m] * kvw, k.data() + (size_t)m * kvw, (size_t)kvw * 4);
                memcpy(vc[slot[m]][l].data() + (size_t)pos[m] * kvw, v.data() + (size_t)m * kvw, (size_t)kvw * 4);
            }
            lap(t_other);
            attention(l, M);
            dbg("att", l, att_out.data(), qw);
            lap(t_att);
            matmul(L.wo, att_out.data(), qw, M, part.data(), dim);
            dbg("wo_part", l, part.data(), dim);
            lap(t_matmul);
            sync_and_merge(M);
            dbg("x_att", l, x.data(), dim);
            lap(t_sync);
            rmsnorm(x.data(), L.rms_ffn.data(), y.data(), M, dim);
            lap(t_other);
            matmul(L.w1, y.data(), dim, M, hb.data(), hw);
            matmul(L.w3, y.data(), dim, M, hl.data(), hw);
            lap(t_matmul);
            pool->run(M * hw, [&](int, int b, int e) {
                for (int i = b; i < e; i++) hb[i] = activate(hb[i], h.hidden_act) * hl[i];
            });
            lap(t_other);
            dbg("silu_mul", l, hb.data(), hw);
            matmul(L.w2, hb.data(), hw, M, part.data(), dim);
            dbg("w2_part", l, part.data(), dim);
            lap(t_matmul);
            sync_and_merge(M);
            dbg("x_ffn", l, x.data(), dim);
            lap(t_sync);
        }
        rmsnorm(x.data() + (size_t)lf * dim, rms_final.data(), y.data(), M - lf, dim);
        lap(t_other);
        matmul(wcls, y.data(), dim, M - lf, lg.data(), h.vocab / n);
        dbg("logits", h.n_layers, lg.data(), h.vocab / n);
        lap(t_matmul);
    }
};

// ---- wire protocol root <-> worker ----------------------------------------------------------------------------
constexpr int32_t kHello = 23571114;            // same handshake constant as the Python bootstrap (cli/app.py)
enum Tag : int32_t { T_END = 0, T_WQ, T_WK, T_WV, T_WO, T_W1, T_W2, T_W3, T_RMS_ATT, T_RMS_FFN, T_RMS_FINAL, T_WCLS };

void send_mat(int fd, int tag, int layer, const Mat& m, NetStats* st) {
    const int64_t nb = m.q40 ? (int64_t)m.q.size() : (int64_t)m.f.size() * 4;
    int32_t hd[5] = {tag, layer, m.d, m.n, m.q40 ? 1 : 0};
    send_all(fd, hd, sizeof hd, st);
    send_all(fd, &nb, 8, st);
    send_all(fd, m.q40 ? (const void*)m.q.data() : (const void*)m.f.data(), (size_t)nb, st);
}

void send_vec(int fd, int tag, int layer, const std::vector<float>& v, NetStats* st) {
    const int64_t nb = (int64_t)v.size() * 4;
    int32_t hd[5] = {tag, layer, 1, (int32_t)v.size(), 0};
    send_all(fd, hd, sizeof hd, st);
    send_all(fd, &nb, 8, st);
    send_all(fd, v.data(), (size_t)nb, st);
}

// slices of node `r` cut out of the mapped file (reference: loadLlmNetWeight + NnRootWeightLoader, llm.cpp:447-483)
struct FileWalker {
    const MappedFile& f;
    const Header& h;
    FileWalker(const MappedFile& f_, const Header& h_) : f(f_), h(h_) {}
    size_t off_embedding() const { return (size_t)h.header_size; }
    size_t layer_bytes() const {
        const int wt = h.weight_type;
        return (size_t)h.dim * row_bytes(wt, h.dim) * 2 + (size_t)h.kv_dim() * row_bytes(wt, h.dim) * 2
               + (size_t)h.hidden_dim * row_bytes(wt, h.dim) * 2 + (size_t)h.dim * row_bytes(wt, h.hidden_dim) + (size_t)h.dim * 8;
    }
    void check_size() const {
        const size_t want = off_embedding() + (size_t)h.vocab * h.dim * 4 + layer_bytes() * h.n_layers + (size_t)h.dim * 4
                            + (size_t)h.vocab * row_bytes(h.weight_type, h.dim);
        if (want != f.size) throw std::runtime_error("Model file size does not match its header (" + std::to_string(f.size) +
                                                     " bytes, expected " + std::to_string(want) + ")");
    }
    void layer(int l, int r, int n, LayerW& L) const {
        const int wt = h.weight_type, dim = h.dim, kv = h.kv_dim(), hid = h.hidden_dim;
        const uint8_t* p = f.data + off_embedding() + (size_t)h.vocab * dim * 4 + layer_bytes() * l;
        slice_rows(p, wt, dim, r * (dim / n), dim / n, L.wq);  p += (size_t)dim * row_bytes(wt, dim);
        slice_rows(p, wt, dim, r * (kv / n), kv / n, L.wk);    p += (size_t)kv * row_bytes(wt, dim);
        slice_rows(p, wt, dim, r * (kv / n), kv / n, L.wv);    p += (size_t)kv * row_bytes(wt, dim);
        slice_cols(p, wt, dim, dim, r * (dim / n), dim / n, L.wo);  p += (size_t)dim * row_bytes(wt, dim);
        slice_rows(p, wt, dim, r * (hid / n), hid / n, L.w1);  p += (size_t)hid * row_bytes(wt, dim);
        slice_cols(p, wt, hid, dim, r * (hid / n), hid / n, L.w2);  p += (size_t)dim * row_bytes(wt, hid);
        slice_rows(p, wt, dim, r * (hid / n), hid / n, L.w3);  p += (size_t)hid * row_bytes(wt, dim);
        L.rms_att.resize(dim);
        L.rms_ffn.resize(dim);
        memcpy(L.rms_att.data(), p, (size_t)dim * 4);
        memcpy(L.rms_ffn.data(), p + (size_t)dim * 4, (size_t)dim * 4);
    }
    const uint8_t* tail() const { return f.data + off_embedding() + (size_t)h.vocab * h.dim * 4 + layer_bytes() * h.n_layers; }
};

void validate_slicing(const Header& h, int n) {         // (reference: app.cpp:237-240 and the slicers' asserts)
    if (n < 1 || (n & (n - 1))) throw std::runtime_error("The number of nodes must be a power of two");
    if (n > h.n_kv_heads)
        throw std::runtime_error("This version does not support more nodes than the number of KV heads in the model");
    if (h.dim % n || h.hidden_dim % n || h.vocab % n || h.n_heads % n || h.n_kv_heads % n) throw std::runtime_error("Model dimensions are not divisible by the number of nodes");
    if (h.weight_type == F_Q40 && ((h.dim / n) % QK || (h.hidden_dim / n) % QK))
        throw std::runtime_error("Q40 slices must be multiples of 32 elements");
    if (h.weight_type != F_Q40 && h.weight_type != F_32 && h.weight_type != F_16) throw std::runtime_error("Unsupported weight type");
}

struct Root {
    Node node;
    std::vector<float> logits_part;             // gather buffer for the workers' slices
    std::string error;
};

void set_error(char* err, int errlen, const std::string& m) {
    if (err && errlen > 0) snprintf(err, (size_t)errlen, "%s", m.c_str());
}

}  // namespace

// =====================================================================================================
// C ABI
// =====================================================================================================
extern "C" {

// Root node: maps the model, keeps slice 0, ships every worker its configuration and weight slices (workers need
// neither the model file nor any knowledge of the model), builds the mesh.  sync_q80: 1 = Q80 sync pipe, 0 = f32.
void* dl_cpu_root_new(const char* model_path, int n_workers, const char** hosts, const int* ports, int n_threads,
                      int max_rows, int n_slots, int max_seq_len, int sync_q80, char* err, int errlen) {
    auto* R = new Root();
    try {
        MappedFile f;
        if (!f.open(model_path)) throw std::runtime_error(std::string("Cannot open model file ") + model_path);
        Node& nd = R->node;
        load_header(f, max_seq_len, nd.c.h);
        const Header& h = nd.c.h;
        const int n = n_workers + 1;
        validate_slicing(h, n);
        if (h.weight_type == F_Q40 && !sync_q80) throw std::runtime_error("This version supports only Q40 weights with Q80 sync type");
        FileWalker fw(f, h);
        fw.check_size();
        nd.c.n_nodes = n;
        nd.c.node = 0;
        nd.c.max_rows = max_rows;
        nd.c.n_slots = n_slots;
        nd.c.sync_q80 = sync_q80 ? 1 : 0;
        nd.c.n_threads = std::max(1, n_threads);
        nd.mesh.n_nodes = n;
        nd.mesh.node = 0;
        nd.mesh.peer.assign(n, -1);
        // connect, hand out node index + peer list + configuration (reference: NnNetwork::connect + NnRootConfigWriter)
        for (int w = 0; w < n_workers; w++) {
            const int fd = connect_to(hosts[w], ports[w], 30);
            nd.mesh.peer[w + 1] = fd;
            NodeConfig wc = nd.c;
            wc.node = w + 1;
            int32_t hello[2] = {kHello, n_workers};
            send_all(fd, hello, sizeof hello, &nd.mesh.stats);
            send_all(fd, &wc, sizeof wc, &nd.mesh.stats);
            for (int o = 0; o < n_workers; o++) {
                char hostbuf[256] = {0};
                snprintf(hostbuf, sizeof hostbuf, "%s", hosts[o]);
                int32_t port = ports[o];
                send_all(fd, hostbuf, sizeof hostbuf, &nd.mesh.stats);
                send_all(fd, &port, 4, &nd.mesh.stats);
            }
            int32_t ack = 0;
            recv_all(fd, &ack, 4, &nd.mesh.stats);
            if (ack != kHello) throw std::runtime_error("worker refused the configuration");
        }
        // weights: own slices stay, the others are streamed (reference: NnRootWeightLoader, nn-network.cpp:766-854)
        nd.embedding.resize((size_t)h.vocab * h.dim);
        memcpy(nd.embedding.data(), f.data + fw.off_embedding(), nd.embedding.size() * 4);
        nd.layers.resize(h.n_layers);
        for (int l = 0; l < h.n_layers; l++) {
            fw.layer(l, 0, n, nd.layers[l]);
            for (int w = 1; w < n; w++) {
                LayerW L;
                fw.layer(l, w, n, L);
                const int fd = nd.mesh.peer[w];
                NetStats* st = &nd.mesh.stats;
                send_mat(fd, T_WQ, l, L.wq, st); send_mat(fd, T_WK, l, L.wk, st); send_mat(fd, T_WV, l, L.wv, st);
                send_mat(fd, T_WO, l, L.wo, st); send_mat(fd, T_W1, l, L.w1, st); send_mat(fd, T_W2, l, L.w2, st);
                send_mat(fd, T_W3, l, L.w3, st);
                send_vec(fd, T_RMS_ATT, l, L.rms_att, st); send_vec(fd, T_RMS_FFN, l, L.rms_ffn, st);
            }
        }
        const uint8_t* tail = fw.tail();
        nd.rms_final.resize(h.dim);
        memcpy(nd.rms_final.data(), tail, (size_t)h.dim * 4);
        const uint8_t* cls = tail + (size_t)h.dim * 4;
        slice_rows(cls, h.weight_type, h.dim, 0, h.vocab / n, nd.wcls);
        for (int w = 1; w < n; w++) {
            Mat m;
            slice_rows(cls, h.weight_type, h.dim, w * (h.vocab / n), h.vocab / n, m);
            const int fd = nd.mesh.peer[w];
            send_vec(fd, T_RMS_FINAL, -1, nd.rms_final, &nd.mesh.stats);
            send_mat(fd, T_WCLS, -1, m, &nd.mesh.stats);
            int32_t end[5] = {T_END, 0, 0, 0, 0};
            int64_t zero = 0;
            send_all(fd, end, sizeof end, &nd.mesh.stats);
            send_all(fd, &zero, 8, &nd.mesh.stats);
        }
        for (int w = 1; w < n; w++) {               // every worker has loaded its slices and joined the mesh
            int32_t ack = 0;
            recv_all(nd.mesh.peer[w], &ack, 4, &nd.mesh.stats);
            if (ack != kHello) throw std::runtime_error("worker failed to load its weights");
        }
        nd.init_buffers();
        R->logits_part.resize((size_t)max_rows * h.vocab / n);
        return R;
    } catch (const std::exception& e) {
        set_error(err, errlen, e.what());
        delete R;
        return nullptr;
    }
}

// Runs M rows (token, position, KV slot each) through all nodes; logits of rows [logits_from, M) -> logits_out
// ([M - logits_from][vocab], f32).  0 on success; -1 + dl_cpu_last_error on failure.
int dl_cpu_forward(void* hp, const int* tokens, const int* pos, const int* slots, int M, int logits_from, float* logits_out) {
    auto* R = (Root*)hp;
    Node& nd = R->node;
    try {
        const Header& h = nd.c.h;
        const int n = nd.N(), dim = h.dim;
        if (M < 1 || M > nd.c.max_rows || logits_from < 0 || logits_from >= M) throw std::runtime_error("bad batch shape");
        for (int m = 0; m < M; m++) {
            if (tokens[m] < 0 || tokens[m] >= h.vocab) throw std::runtime_error("token id outside the vocabulary");
            nd.pos[m] = pos[m];
            nd.slot[m] = slots[m];
            memcpy(nd.x.data() + (size_t)m * dim, nd.embedding.data() + (size_t)tokens[m] * dim, (size_t)dim * 4);   // OP_EMBEDDING
        }
        // S5 control packet + S1 broadcast of x (reference: app.cpp:198-201, llm.cpp:193) in one message per worker
        for (int w = 1; w < n; w++) {
            const int fd = nd.mesh.peer[w];
            int32_t ctl[2] = {M, logits_from};
            send_all(fd, ctl, sizeof ctl, &nd.mesh.stats);
            send_all(fd, nd.pos.data(), (size_t)M * 4, &nd.mesh.stats);
            send_all(fd, nd.slot.data(), (size_t)M * 4, &nd.mesh.stats);
            send_all(fd, nd.x.data(), (size_t)M * dim * 4, &nd.mesh.stats);
        }
        nd.forward_node(M, logits_from);
        // S4: logits slices to the root (reference: SYNC_NODE_SLICES_EXCEPT_ROOT, llm.cpp:432)
        const int rows = M - logits_from, vs = h.vocab / n;
        for (int r = 0; r < rows; r++) memcpy(logits_out + (size_t)r * h.vocab, nd.lg.data() + (size_t)r * vs, (size_t)vs * 4);
        for (int w = 1; w < n; w++) {
            recv_all(nd.mesh.peer[w], R->logits_part.data(), (size_t)rows * vs * 4, &nd.mesh.stats);
            for (int r = 0; r < rows; r++)
                memcpy(logits_out + (size_t)r * h.vocab + (size_t)w * vs, R->logits_part.data() + (size_t)r * vs, (size_t)vs * 4);
        }
        return 0;
    } catch (const std::exception& e) {
        R->error = e.what();
        return -1;
    }
}

const char* dl_cpu_last_error(void* hp) { return ((Root*)hp)->error.c_str(); }

// out[0] = bytes sent, out[1] = bytes received, out[2] = microseconds spent in node synchronisation
// (reference: NnNetwork::getStats + the executor's sync timer: the "Sync ms | Sent kB Recv kB" of dllama inference)
void dl_cpu_stats(void* hp, long long* out, int reset) {
    Node& nd = ((Root*)hp)->node;
    out[0] = nd.mesh.stats.sent;
    out[1] = nd.mesh.stats.recv;
    out[2] = (long long)(nd.mesh.sync_s * 1e6);
    if (reset) { nd.mesh.stats.sent = 0; nd.mesh.stats.recv = 0; nd.mesh.sync_s = 0.0; }
}

void dl_cpu_free(void* hp) {
    auto* R = (Root*)hp;
    if (!R) return;
    for (int w = 1; w < R->node.N(); w++) {        // M = 0 stops the workers (reference: app.cpp:204-209)
        int32_t ctl[2] = {0, 0};
        try { send_all(R->node.mesh.peer[w], ctl, sizeof ctl, nullptr); } catch (...) {}
    }
    delete R;
}

// `dllama worker`: listen, take one root, run its batches until it stops or disconnects; returns then (0), so the
// caller decides whether to listen again (reference: runWorkerApp, src/app.cpp:405-464).  n_threads <= 0: as the
// root says.  < 0 with a message in err on failure.
int dl_cpu_worker_serve(int port, int n_threads, char* err, int errlen) {
    int lfd = -1;
    try {
        lfd = socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
        int one = 1;
        setsockopt(lfd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
        sockaddr_in a{};
        a.sin_family = AF_INET;
        // DLLAMA_BIND_ADDR: listen on one interface only (default: all, like the reference's worker)
        const char* bind_addr = getenv("DLLAMA_BIND_ADDR");
        a.sin_addr.s_addr = htonl(INADDR_ANY);
        if (bind_addr != nullptr && *bind_addr && inet_pton(AF_INET, bind_addr, &a.sin_addr) != 1)
            throw std::runtime_error(std::string("DLLAMA_BIND_ADDR is not an IPv4 address: ") + bind_addr);
        a.sin_port = htons((uint16_t)port);
        if (bind(lfd, (sockaddr*)&a, sizeof a) != 0 || listen(lfd, 16) != 0)
            throw std::runtime_error(std::string("cannot listen on port ") + std::to_string(port) + ": " + strerror(errno));
        printf("Listening on %s:%d...\n", (bind_addr != nullptr && *bind_addr) ? bind_addr : "0.0.0.0", port);
        fflush(stdout);
        Node nd;
        const int rfd = accept(lfd, nullptr, nullptr);
        if (rfd < 0) throw std::runtime_error("accept failed");
        set_nodelay(rfd);
        int32_t hello[2];
        recv_all(rfd, hello, sizeof hello, nullptr);
        if (hello[0] != kHello) throw std::runtime_error("bad handshake from the root");
        const int n_workers = hello[1];
        if (n_workers < 1 || n_workers > 63) throw std::runtime_error("bad worker count in the handshake");
        recv_all(rfd, &nd.c, sizeof nd.c, &nd.mesh.stats);
        {
            // everything below is sized from this packet: refuse values that cannot describe a model before any buffer
            // is allocated or indexed with them (the peer is trusted like the reference trusts its root, but a wrong
            // packet must end in an error message, not in memory corruption)
            const Header& hh = nd.c.h;
            const bool ok = nd.c.n_nodes == n_workers + 1 && nd.c.node >= 1 && nd.c.node < nd.c.n_nodes &&
                            nd.c.max_rows >= 1 && nd.c.max_rows <= (1 << 16) && nd.c.n_slots >= 1 && nd.c.n_slots <= (1 << 16) &&
                            hh.dim >= 32 && hh.dim <= (1 << 17) && hh.hidden_dim >= 32 && hh.hidden_dim <= (1 << 20) &&
                            hh.n_layers >= 1 && hh.n_layers <= 1024 && hh.n_heads >= 1 && hh.n_heads <= 1024 &&
                            hh.n_kv_heads >= 1 && hh.n_kv_heads <= hh.n_heads && hh.vocab >= 1 && hh.vocab <= (1 << 22) &&
                            hh.seq_len >= 1 && hh.seq_len <= (1 << 22) && hh.dim % hh.n_heads == 0;
            if (!ok) throw std::runtime_error("bad node configuration from the root");
        }
        if (n_threads > 0) nd.c.n_threads = n_threads;
        std::vector<std::string> hosts(n_workers);
        std::vector<int> ports(n_workers);
        for (int o = 0; o < n_workers; o++) {
            char hostbuf[256];
            int32_t p;
            recv_all(rfd, hostbuf, sizeof hostbuf, &nd.mesh.stats);
            recv_all(rfd, &p, 4, &nd.mesh.stats);
            hostbuf[255] = 0;
            hosts[o] = hostbuf;
            ports[o] = p;
        }
        const int n = nd.c.n_nodes, me = nd.c.node;
        validate_slicing(nd.c.h, n);
        nd.mesh.n_nodes = n;
        nd.mesh.node = me;
        nd.mesh.peer.assign(n, -1);
        nd.mesh.peer[0] = rfd;
        send_all(rfd, &kHello, 4, &nd.mesh.stats);
        printf("⭕ The root node has connected, node %d/%d\n", me, n);
        // weight slices (reference: NnWorkerWeightReader::read, nn-network.cpp:876-901)
        nd.layers.resize(nd.c.h.n_layers);
        long long wbytes = 0;
        while (true) {
            int32_t hd[5];
            int64_t nb;
            recv_all(rfd, hd, sizeof hd, &nd.mesh.stats);
            recv_all(rfd, &nb, 8, &nd.mesh.stats);
            if (hd[0] == T_END) break;
            const int tag = hd[0], l = hd[1];
            const bool per_layer = tag >= T_WQ && tag <= T_RMS_FFN;
            if (nb < 0 || (per_layer && (l < 0 || l >= nd.c.h.n_layers))) throw std::runtime_error("malformed weight record");
            auto read_mat = [&](Mat& m) {
                m.d = hd[2]; m.n = hd[3]; m.q40 = hd[4] != 0;
                // the slice this node owns is fixed by the header and the node count: anything else is a corrupt record
                const Header& hh = nd.c.h;
                const int qd = hh.dim / n, kvd = hh.kv_dim() / n, hid = hh.hidden_dim / n;
                int ed = 0, en = 0;
                switch (tag) {
                    case T_WQ: ed = qd; en = hh.dim; break;
                    case T_WK: case T_WV: ed = kvd; en = hh.dim; break;
                    case T_WO: ed = hh.dim; en = qd; break;
                    case T_W1: case T_W3: ed = hid; en = hh.dim; break;
                    case T_W2: ed = hh.dim; en = hid; break;
                    default: ed = hh.vocab / n; en = hh.dim; break;      // T_WCLS
                }
                const long long want = m.q40 ? (long long)ed * (en / QK) * Q40_BYTES : (long long)ed * en * 4;
                if (m.d != ed || m.n != en || nb != want) throw std::runtime_error("weight record does not match the model header");
                if (m.q40) { m.q.resize((size_t)nb); recv_all(rfd, m.q.data(), (size_t)nb, &nd.mesh.stats); }
                else { m.f.resize((size_t)nb / 4); recv_all(rfd, m.f.data(), (size_t)nb, &nd.mesh.stats); }
            };
            auto read_vec = [&](std::vector<float>& v) {
                if (nb != (long long)nd.c.h.dim * 4) throw std::runtime_error("norm record does not match the model header");
                v.resize((size_t)nb / 4);
                recv_all(rfd, v.data(), (size_t)nb, &nd.mesh.stats);
            };
            switch (tag) {
                case T_WQ: read_mat(nd.layers[l].wq); break;
                case T_WK: read_mat(nd.layers[l].wk); break;
                case T_WV: read_mat(nd.layers[l].wv); break;
                case T_WO: read_mat(nd.layers[l].wo); break;
                case T_W1: read_mat(nd.layers[l].w1); break;
                case T_W2: read_mat(nd.layers[l].w2); break;
                case T_W3: read_mat(nd.layers[l].w3); break;
                case T_RMS_ATT: read_vec(nd.layers[l].rms_att); break;
                case T_RMS_FFN: read_vec(nd.layers[l].rms_ffn); break;
                case T_RMS_FINAL: read_vec(nd.rms_final); break;
                case T_WCLS: read_mat(nd.wcls); break;
                default: throw std::runtime_error("unknown weight record");
            }
            wbytes += nb;
        }
        printf("💿 Weights loaded (%lld kB)\n", wbytes / 1024);
        // mesh between workers: connect to higher indices, accept from lower ones (reference: nn-network.cpp:264-319)
        for (int o = me + 1; o < n; o++) {
            const int fd = connect_to(hosts[o - 1], ports[o - 1], 30);
            int32_t id[2] = {kHello, me};
            send_all(fd, id, sizeof id, &nd.mesh.stats);
            nd.mesh.peer[o] = fd;
        }
        for (int k = 1; k < me; k++) {
            const int fd = accept(lfd, nullptr, nullptr);
            if (fd < 0) throw std::runtime_error("accept failed");
            set_nodelay(fd);
            int32_t id[2];
            recv_all(fd, id, sizeof id, &nd.mesh.stats);
            if (id[0] != kHello || id[1] < 1 || id[1] >= me || nd.mesh.peer[id[1]] >= 0) throw std::runtime_error("bad peer handshake");
            nd.mesh.peer[id[1]] = fd;
        }
        close(lfd);
        lfd = -1;
        nd.init_buffers();
        send_all(rfd, &kHello, 4, &nd.mesh.stats);
        printf("🚁 Worker is ready (%d threads)\n", nd.pool->size());
        fflush(stdout);
        const int dim = nd.c.h.dim;
        while (true) {
            int32_t ctl[2];
            try {
                recv_all(rfd, ctl, sizeof ctl, &nd.mesh.stats);
            } catch (const std::exception&) {
                break;                                  // root went away: back to the caller, which listens again
            }
            const int M = ctl[0], lf = ctl[1];
            if (M <= 0) break;
            if (M > nd.c.max_rows || lf < 0 || lf >= M) throw std::runtime_error("bad control packet");
            try {
                recv_all(rfd, nd.pos.data(), (size_t)M * 4, &nd.mesh.stats);
                recv_all(rfd, nd.slot.data(), (size_t)M * 4, &nd.mesh.stats);
                recv_all(rfd, nd.x.data(), (size_t)M * dim * 4, &nd.mesh.stats);
                nd.forward_node(M, lf);
                send_all(rfd, nd.lg.data(), (size_t)(M - lf) * (nd.c.h.vocab / n) * 4, &nd.mesh.stats);
            } catch (const std::exception& e) {
                // a root (or peer) that dies in the middle of a batch: drop everything and wait for the next root
                // (reference: the worker catches Nn{Read,Write}NetworkException and re-serves, src/app.cpp:455-461)
                printf("🚨 Network is closed: %s\n", e.what());
                break;
            }
        }
        printf("🚁 Root disconnected\n");
        fflush(stdout);
        return 0;
    } catch (const std::exception& e) {
        if (lfd >= 0) close(lfd);
        set_error(err, errlen, e.what());
        return -1;
    }
}

// ---- codec entry points (tests, converters; reference: quantizeF32toQ80 / dequantizeQ80toF32 / dequantizeQ40toF32) ----
void dl_cpu_quantize_q80(const float* x, uint8_t* out, int n, int force_scalar) {
    if (force_scalar) quantize_q80_scalar(x, out, n); else quantize_q80(x, out, n);
}
void dl_cpu_dequantize_q80(const uint8_t* in, float* y, int n) { dequantize_q80(in, y, n); }
void dl_cpu_dequantize_q40(const uint8_t* in, float* y, int n) {
    for (int b = 0; b < n / QK; b++) {
        const uint8_t* p = in + (size_t)b * Q40_BYTES;
        uint16_t dh;
        memcpy(&dh, p, 2);
        const float d = f16_to_f32(dh);
        for (int j = 0; j < 16; j++) {
            y[b * QK + j] = (float)((int)(p[2 + j] & 0x0f) - 8) * d;
            y[b * QK + j + 16] = (float)((int)(p[2 + j] >> 4) - 8) * d;
        }
    }
}
// Q40 quantiser of the converter (reference: writer.py:29-53 / quantizeF32toQ40, nn-quants.cpp:193-227)
void dl_cpu_quantize_q40(const float* x, uint8_t* out, int n) {
    for (int b = 0; b < n / QK; b++) {
        const float* g = x + b * QK;
        float amax = 0.f, mx = 0.f;
        for (int j = 0; j < QK; j++) if (fabsf(g[j]) > amax) { amax = fabsf(g[j]); mx = g[j]; }
        const float d = mx / -8.f;
        const float id = d != 0.f ? 1.f / d : 0.f;
        uint8_t* o = out + (size_t)b * Q40_BYTES;
        const uint16_t dh = f32_to_f16(d);
        memcpy(o, &dh, 2);
        for (int j = 0; j < 16; j++) {
            const int a = std::min(15, (int)(g[j] * id + 8.5f)), c = std::min(15, (int)(g[j + 16] * id + 8.5f));
            o[2 + j] = (uint8_t)(std::max(0, a) | (std::max(0, c) << 4));
        }
    }
}
unsigned short dl_cpu_f32_to_f16(float f) { return f32_to_f16(f); }
float dl_cpu_f16_to_f32(unsigned short h) { return f16_to_f32(h); }
// y[m][r] for rows [0, d): scalar or dispatched kernel (tests compare the two)
// which: 0 = best available (what the engine uses), 1 = scalar, 2 = AVX2 (falls back to scalar where unsupported)
void dl_cpu_matmul_q40_q80(const uint8_t* w, int d, int n, const uint8_t* xq, int M, float* y, int which) {
    if (which == 1) matmul_q40_q80_scalar(w, 0, d, n / QK, xq, M, y, (size_t)d);
#ifdef DL_X86
    else if (which == 2 && has_avx2()) matmul_q40_q80_avx2(w, 0, d, n / QK, xq, M, y, (size_t)d);
#endif
    else if (which == 2) matmul_q40_q80_scalar(w, 0, d, n / QK, xq, M, y, (size_t)d);
    else matmul_q40_q80(w, 0, d, n / QK, xq, M, y, (size_t)d);
}
int dl_cpu_has_avx2(void) { return has_avx2() ? 1 : 0; }
int dl_cpu_has_vnni(void) { return has_vnni() ? 1 : 0; }

// ---- op-level entry points (unit tests against the reference's goldens, src/nn/nn-cpu-ops-test.cpp) ----
float dl_cpu_inv_rms(const float* x, int n, float eps) { return inv_rms(x, n, eps); }
void dl_cpu_rms_norm(const float* x, const float* w, float* y, int n, float eps) { rms_norm_row(x, w, inv_rms(x, n, eps), y, n); }
void dl_cpu_softmax(float* x, int n) { softmax_inplace(x, n); }
void dl_cpu_activation(float* x, int n, int act /*1 = SiLU, 0 = GELU*/) { for (int i = 0; i < n; i++) x[i] = activate(x[i], act); }
void dl_cpu_rope(float* row, int width, int head_size, int pos, float theta, int rope_type, float factor, float low,
                 float high, int orig_len) {
    std::vector<float> freq(head_size / 2);
    for (int j = 0; j < head_size / 2; j++) freq[j] = rope_frequency(j, head_size, theta, rope_type, factor, low, high, orig_len);
    rope_row(row, width, head_size, pos, freq.data());
}

}  // extern "C"
