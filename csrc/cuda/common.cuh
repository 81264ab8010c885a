// Shared host/device helpers for the sm_100a library.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

#include "b200cnn.h"

namespace b200cnn {

// Records the message returned by b200cnn_last_error(); returns `code` for `return set_error(...)`.
int set_error(int code, const char* fmt, ...);

inline cudaStream_t as_stream(b200cnn_stream s) { return static_cast<cudaStream_t>(s); }

// After a kernel launch: pick up launch-configuration errors without synchronising.
#define B200CNN_CHECK_LAUNCH(name)                                                              \
    do {                                                                                        \
        cudaError_t e__ = cudaGetLastError();                                                   \
        if (e__ != cudaSuccess) return ::b200cnn::set_error(int(e__), "%s: %s", name, cudaGetErrorString(e__)); \
    } while (0)

#define B200CNN_CUDA(call)                                                                      \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess)                                                                 \
            return ::b200cnn::set_error(int(e__), "%s failed: %s", #call, cudaGetErrorString(e__)); \
    } while (0)

constexpr int kInvalidValue = 1;  // cudaErrorInvalidValue

// ---- programmatic dependent launch (PDL) -----------------------------------------------------------
// Every kernel of the training step is launched with programmatic stream serialization: the next
// kernel in the stream may be scheduled while this one still runs, so its prologue (barrier init, TMEM
// allocation, tensor-map prefetch, launch latency itself) overlaps our body.  Contract: a kernel calls
// pdl_wait() before its first global-memory access (blocks until the previous kernel has completed and
// its writes are visible) and pdl_launch_dependents() right after, which lets the NEXT kernel start its
// own prologue.  At most two kernels of a stream are ever in flight.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();  // B200CNN_PDL=0 in the environment turns the launch attribute off

// Kernels launched on the "critical" stream of the calling thread get the highest launch priority the device offers:
// the training step forks its weight-gradient kernels onto a side stream, and when one of those still has CTAs waiting
// for an SM (grids of 2+ waves) the block scheduler would otherwise finish dispatching it before it starts the kernel
// that continues the critical path.  (Captured into CUDA-graph kernel nodes like any other launch attribute.)
// B200CNN_PRIORITY=0 turns it off.
struct CriticalStreamScope {
    explicit CriticalStreamScope(cudaStream_t st);
    ~CriticalStreamScope();
    cudaStream_t saved;
};
cudaStream_t critical_stream();  // of this thread; nullptr = none
int critical_priority();         // 0 = none / disabled

// cluster_z > 1: thread-block clusters of (1, 1, cluster_z) CTAs (grid.z must be a multiple of it).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                         int cluster_z, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[3];
    int n = 0;
    if (st != nullptr && st == critical_stream() && critical_priority() != 0) {
        attr[n].id = cudaLaunchAttributePriority;
        attr[n].val.priority = critical_priority();
        n++;
    }
    if (pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        n++;
    }
    if (cluster_z > 1) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = 1;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = unsigned(cluster_z);
        n++;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    return launch_kernel_cluster(kernel, grid, block, smem, st, 1, args...);
}

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Derivative of an activation expressed through its OUTPUT value, like the reference's
// tanh_g / relu_g (cnn.c:51-57).
__device__ __forceinline__ float act_grad_from_output(int act, float y) {
    if (act == B200CNN_ACT_TANH) return 1.0f - y * y;
    if (act == B200CNN_ACT_RELU) return y > 0.0f ? 1.0f : 0.0f;
    return 1.0f;
}

__device__ __forceinline__ float apply_act(int act, float v) {
    if (act == B200CNN_ACT_TANH) return tanhf(v);
    if (act == B200CNN_ACT_RELU) return v > 0.0f ? v : 0.0f;
    return v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace b200cnn
