// Does cp.async.bulk.prefetch.L2 shorten a later cp.async.bulk load?  One CTA per SM; each CTA times (SM clock) the
// arrival of a 9216-byte bulk copy from (a) cold memory, (b) memory it prefetched to L2 a few microseconds earlier,
// (c) memory it just loaded (L2 hit by construction).  nvcc -arch=sm_100a -O3 l2_prefetch_probe.cu -o probe && ./probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(const uint8_t* buf, size_t stride, long long* out, int mode) {
    extern __shared__ __align__(128) uint8_t sm[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(sm);
    uint8_t* dst = sm + 128;
    const uint8_t* src = buf + (size_t)blockIdx.x * stride;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t parity = 0;
        for (int rep = 0; rep < 8; rep++) {
            const uint8_t* p = src + (size_t)rep * 9216 * 4;
            if (mode == 1) {
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(9216) : "memory");
                long long t0 = clock64();
                while (clock64() - t0 < 10000) {}
            } else if (mode == 2) {      // load it once (brings it to L2), then time the second load
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(9216) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst)), "l"(p), "r"(9216), "r"(s32(bar)) : "memory");
                uint32_t ok = 0;
                while (!ok) asm volatile("{ .reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2; selp.u32 %0, 1, 0, q; }" : "=r"(ok) : "r"(s32(bar)), "r"(parity) : "memory");
                parity ^= 1;
            }
            long long t0 = clock64();
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(9216) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst)), "l"(p), "r"(9216), "r"(s32(bar)) : "memory");
            uint32_t ok = 0;
            while (!ok) asm volatile("{ .reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2; selp.u32 %0, 1, 0, q; }" : "=r"(ok) : "r"(s32(bar)), "r"(parity) : "memory");
            parity ^= 1;
            out[blockIdx.x * 8 + rep] = clock64() - t0;
        }
    }
}
int main() {
    const int G = 148;
    const size_t stride = 1 << 20;
    uint8_t* buf; long long* out;
    cudaMalloc(&buf, (size_t)G * stride * 4);
    cudaMalloc(&out, G * 8 * sizeof(long long));
    uint8_t* flush; cudaMalloc(&flush, 512u << 20);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
    const char* names[3] = {"cold HBM", "after bulk prefetch to L2 (+10k clk)", "second load (L2 hit)"};
    for (int mode = 0; mode < 3; mode++) {
        cudaMemset(flush, mode, 512u << 20);        // push everything out of L2
        cudaDeviceSynchronize();
        probe<<<G, 32, 16384>>>(buf + (size_t)mode * G * stride, stride, out, mode);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        static long long h[G * 8];
        cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
        double sum = 0; long long mn = 1ll << 60, mx = 0;
        for (int i = 0; i < G * 8; i++) { sum += h[i]; if (h[i] < mn) mn = h[i]; if (h[i] > mx) mx = h[i]; }
        printf("%-40s mean %.0f clk  min %lld  max %lld\n", names[mode], sum / (G * 8), mn, mx);
    }
    return 0;
}
