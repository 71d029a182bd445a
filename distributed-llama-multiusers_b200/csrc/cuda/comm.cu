// Peer-memory plumbing: symmetric buffers shared between the one-process-per-GPU ranks through
// CUDA IPC (NVLink / NVSwitch peer mappings), plus a flag barrier kernel.
//
// This replaces the reference's TCP socket layer for the data plane (NnNetwork::writeMany /
// readMany busy loops, src/nn/nn-network.cpp:419-491): after the handles are exchanged at
// start-up, all traffic is ordinary loads/stores issued from inside the compute kernels
// (epilogue.cuh DL_EPI_STORE pushes, prep.cu reduce_norm_quant flag waits).
#include <stdio.h>
#include "common.cuh"
#include "../../../include/dllama_b200.h"

extern void dl_set_error(const char* fmt, ...);

extern "C" int dl_ipc_alloc(void** dev_ptr, long long bytes, uint8_t* handle_out) {
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, (size_t)bytes);
    if (e != cudaSuccess) { dl_set_error("dl_ipc_alloc: cudaMalloc(%lld): %s", bytes, cudaGetErrorString(e)); return -(int)e; }
    e = cudaMemset(p, 0, (size_t)bytes);
    if (e != cudaSuccess) { dl_set_error("dl_ipc_alloc: memset: %s", cudaGetErrorString(e)); return -(int)e; }
    if (handle_out != nullptr) {
        cudaIpcMemHandle_t h;
        e = cudaIpcGetMemHandle(&h, p);
        if (e != cudaSuccess) { dl_set_error("dl_ipc_alloc: cudaIpcGetMemHandle: %s", cudaGetErrorString(e)); cudaFree(p); return -(int)e; }
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
        memcpy(handle_out, &h, 64);
    }
    cudaDeviceSynchronize();
    *dev_ptr = p;
    return 0;
}

extern "C" int dl_ipc_open(const uint8_t* handle, void** dev_ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { dl_set_error("dl_ipc_open: %s", cudaGetErrorString(e)); return -(int)e; }
    *dev_ptr = p;
    return 0;
}

extern "C" int dl_ipc_close(void* dev_ptr) {
    cudaError_t e = cudaIpcCloseMemHandle(dev_ptr);
    if (e != cudaSuccess) { dl_set_error("dl_ipc_close: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}

extern "C" int dl_ipc_free(void* dev_ptr) {
    cudaError_t e = cudaFree(dev_ptr);
    if (e != cudaSuccess) { dl_set_error("dl_ipc_free: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}

struct BarrierArgs {
    uint32_t* peer_flags[DL_MAX_PEERS];   // peer p's flag array (we write slot [rank])
    uint32_t* my_flags;                   // our flag array (peers write slot [their rank])
    int rank, world;
    uint32_t epoch;
};

__global__ void peer_barrier_kernel(const __grid_constant__ BarrierArgs a) {
    const int t = threadIdx.x;
    if (t < a.world) {
        __threadfence_system();
        st_release_sys(a.peer_flags[t] + a.rank, a.epoch);
        wait_flag_ge(a.my_flags + t, a.epoch);
    }
}

extern "C" int dl_peer_barrier(uint32_t* const* flags_on_peers, uint32_t* my_flags, int rank, int world, uint32_t epoch,
                               void* stream) {
    if (world > DL_MAX_PEERS) { dl_set_error("dl_peer_barrier: world > 8"); return -1; }
    BarrierArgs a;
    for (int i = 0; i < world; i++) a.peer_flags[i] = flags_on_peers[i];
    a.my_flags = my_flags;
    a.rank = rank;
    a.world = world;
    a.epoch = epoch;
    peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { dl_set_error("dl_peer_barrier: %s", cudaGetErrorString(e)); return -(int)e; }
    return 0;
}
