// F32 x F32 -> F32 matmul for the f32-weights / f32-buffer path (reference: matmul_F32_F32_F32,
// src/nn/nn-cpu-ops.cpp:182-220): out[m][r] = sum_k W[r][k] * x[m][k], plain f32 FMAs (no tf32, no
// reduced-precision operand), optional accumulate into out (residual / merge-add of a partial sum).
//
// Memory-bound by the weight stream (4 B per weight, read exactly once per launch for M <= 8 rows):
// one warp owns F32_RPW consecutive weight rows, lanes stride over K with 16-byte loads, the (small)
// activation block is re-read through L1/L2, and a butterfly reduction finishes each row.  Wider
// batches run as ceil(M / 8) passes over the weights (prefill of f32 models is not a tuned path).
#include "common.cuh"

#define F32_RPW 2          // weight rows per warp
#define F32_WARPS 8        // warps per CTA
#define F32_MT 8           // activation rows per pass

template <int MT>
__global__ void __launch_bounds__(F32_WARPS * 32) gemm_f32_kernel(const float* __restrict__ w, int d, int n, long long ldw,
                                                                  const float* __restrict__ x, int ldx, float* __restrict__ y,
                                                                  int ldy, int accumulate, int mt /* valid rows <= MT */) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r0 = (blockIdx.x * F32_WARPS + warp) * F32_RPW;
    if (r0 >= d) return;
    float acc[F32_RPW][MT];
#pragma unroll
    for (int i = 0; i < F32_RPW; i++)
#pragma unroll
        for (int m = 0; m < MT; m++) acc[i][m] = 0.0f;
    const int n4 = n >> 2;      // n % 4 == 0 (checked by the host)
    for (int k4 = lane; k4 < n4; k4 += 32) {
        float4 wv[F32_RPW];
#pragma unroll
        for (int i = 0; i < F32_RPW; i++) {
            const int r = min(r0 + i, d - 1);
            wv[i] = __ldcs(reinterpret_cast<const float4*>(w + (size_t)r * ldw) + k4);      // streamed once: evict first
        }
#pragma unroll
        for (int m = 0; m < MT; m++) {
            const float4 xv = __ldg(reinterpret_cast<const float4*>(x + (size_t)min(m, mt - 1) * ldx) + k4);      // rows >= mt: clamped, never stored
#pragma unroll
            for (int i = 0; i < F32_RPW; i++) {
                acc[i][m] = fmaf(wv[i].x, xv.x, acc[i][m]);
                acc[i][m] = fmaf(wv[i].y, xv.y, acc[i][m]);
                acc[i][m] = fmaf(wv[i].z, xv.z, acc[i][m]);
                acc[i][m] = fmaf(wv[i].w, xv.w, acc[i][m]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < F32_RPW; i++)
#pragma unroll
        for (int m = 0; m < MT; m++) {
            const float v = warp_sum(acc[i][m]);
            if (lane == 0 && r0 + i < d && m < mt) {
                float* o = y + (size_t)m * ldy + r0 + i;
                *o = accumulate ? (*o + v) : v;
            }
        }
}

extern void dl_set_error(const char* fmt, ...);

template <int MT>
static cudaError_t launch_f32(const float* w, int d, int n, long long ldw, const float* x, int ldx, float* y, int ldy,
                              int accumulate, int mt, cudaStream_t st) {
    const int rows_per_cta = F32_WARPS * F32_RPW;
    gemm_f32_kernel<MT><<<(d + rows_per_cta - 1) / rows_per_cta, F32_WARPS * 32, 0, st>>>(w, d, n, ldw, x, ldx, y, ldy, accumulate, mt);
    return cudaGetLastError();
}

extern "C" int dl_gemm_f32(const float* w, int d, int n, long long ldw, const float* x, int M, int ldx, float* y, int ldy,
                           int accumulate, void* stream) {
    if (n % 4 != 0 || ldw % 4 != 0 || ldx % 4 != 0) { dl_set_error("dl_gemm_f32: n, ldw, ldx must be multiples of 4"); return -1; }
    if ((((uintptr_t)w) | ((uintptr_t)x)) & 15) { dl_set_error("dl_gemm_f32: w and x must be 16-byte aligned"); return -1; }
    if (M < 1 || d < 1) { dl_set_error("dl_gemm_f32: empty problem"); return -1; }
    cudaStream_t st = (cudaStream_t)stream;
    for (int m0 = 0; m0 < M; m0 += F32_MT) {
        const int mt = M - m0 < F32_MT ? M - m0 : F32_MT;
        const float* xp = x + (size_t)m0 * ldx;
        float* yp = y + (size_t)m0 * ldy;
        cudaError_t e;
        switch (mt) {
            case 1: e = launch_f32<1>(w, d, n, ldw, xp, ldx, yp, ldy, accumulate, mt, st); break;
            case 2: e = launch_f32<2>(w, d, n, ldw, xp, ldx, yp, ldy, accumulate, mt, st); break;
            case 3: case 4: e = launch_f32<4>(w, d, n, ldw, xp, ldx, yp, ldy, accumulate, mt, st); break;
            default: e = launch_f32<8>(w, d, n, ldw, xp, ldx, yp, ldy, accumulate, mt, st); break;
        }
        if (e != cudaSuccess) { dl_set_error("gemm f32 launch: %s", cudaGetErrorString(e)); return -(int)e; }
    }
    return 0;
}
