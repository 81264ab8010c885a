// Peer-mapped memory without any framework: cudaMalloc + CUDA IPC handles.  Used by the native
// multi-process launcher and as the fallback when torch symmetric memory is unavailable.  (The
// reference has no device-side communication at all: its MPI_Allreduce works on host buffers,
// /root/reference/CUDAMPI.c:414.)
#include "../cuda/common.cuh"
#include "b200cnn_comm.h"

using namespace b200cnn;

extern "C" void* b200cnn_ipc_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess || cudaMemset(p, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
        set_error(int(cudaGetLastError()), "ipc_alloc: cannot allocate %zu bytes", bytes);
        return nullptr;
    }
    return p;
}

extern "C" int b200cnn_ipc_free(void* p) {
    B200CNN_CUDA(cudaFree(p));
    return 0;
}

extern "C" int b200cnn_ipc_get_handle(void* p, unsigned char* out64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    cudaIpcMemHandle_t h;
    B200CNN_CUDA(cudaIpcGetMemHandle(&h, p));
    memcpy(out64, &h, 64);
    return 0;
}

extern "C" void* b200cnn_ipc_open(const unsigned char* handle64) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void* p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        set_error(int(e), "ipc_open: %s", cudaGetErrorString(e));
        return nullptr;
    }
    return p;
}

extern "C" int b200cnn_ipc_close(void* p) {
    B200CNN_CUDA(cudaIpcCloseMemHandle(p));
    return 0;
}
