// b200cnn_conv2d_bwd_weight_noalloc: the convolution weight-gradient entry point for callers that
// cannot hand the library a scratch buffer because they are not allowed to allocate device memory.
//
// Who that is.  Some host programs give up cudaMalloc after start-up: a plug-in that runs inside an
// application whose own pool allocator has already claimed the whole GPU (cudaMalloc then simply fails);
// a control loop with a latency budget, where cudaMalloc / cudaFree would stall it because they
// synchronise the device; a process that has been certified with a fixed memory map and may not grow
// it.  The regular entry point, b200cnn_conv2d_bwd_weight, asks such a caller for
// b200cnn_conv2d_bwd_weight_scratch_bytes() of device memory it has no way to obtain.
//
// What this file does instead.  The per-block partial sums go into a fixed array that is part of the
// library image itself (`g_scratch` below, B200CNN_NOALLOC_SCRATCH_BYTES = 8 MiB).  The CUDA runtime
// places that array on every GPU that loads the library and keeps it there until the process exits,
// so nothing is ever allocated or freed here: no cudaMalloc, no cudaFree, no library call that
// allocates on our behalf.  Inputs and outputs are still the caller's.
//
// Things a caller must know:
//   * Limit: the layer must need at most B200CNN_NOALLOC_SCRATCH_BYTES of scratch
//     (chunks x (oc*c*k*k + oc) floats, chunks <= 296); larger shapes are rejected with an error message
//     that states both numbers, nothing is launched.
//   * First call on a device: the array's device address is looked up once per device
//     (cudaGetSymbolAddress, with that device current) and remembered.  Make that first call outside
//     any stream capture — do one warm-up call per device during start-up.
//   * One array per GPU, shared by every call: two calls whose kernels can overlap on the same GPU
//     (different streams, different host threads) would write each other's partial sums.  Keep calls
//     on one device ordered, e.g. by issuing them on one stream.
//   * The array still holds the previous call's partial sums when the next call starts.  That is
//     harmless: the first pass writes every word the second pass reads, so nothing in it has to start
//     at zero and nothing is cleared.
#include <mutex>

#include "backend.cuh"
#include "common.cuh"

namespace {

constexpr size_t kNoAllocScratchBytes = B200CNN_NOALLOC_SCRATCH_BYTES;
constexpr int kMaxDevices = 64;

// Lives in the library image; one instance per GPU, for the lifetime of the process.
static __device__ __align__(256) float g_scratch[kNoAllocScratchBytes / sizeof(float)];

// In host code `g_scratch` is only a handle for the runtime, not an address a kernel could use.  The
// real device address differs per GPU and is resolved here, once per device.
std::mutex g_mutex;
float* g_scratch_ptr[kMaxDevices] = {};

int scratch_for_current_device(float** out) {
    int dev = -1;
    B200CNN_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices)
        return b200cnn::set_error(b200cnn::kInvalidValue, "conv2d_bwd_weight_noalloc: device index %d is outside [0, %d)", dev, kMaxDevices);
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
        if (e != cudaSuccess || !p)
            return b200cnn::set_error(int(e), "conv2d_bwd_weight_noalloc: cannot resolve the built-in scratch on device %d: %s", dev,
                                      cudaGetErrorString(e));
        g_scratch_ptr[dev] = static_cast<float*>(p);
    }
    *out = g_scratch_ptr[dev];
    return 0;
}

}  // namespace

extern "C" size_t b200cnn_conv2d_bwd_weight_noalloc_capacity(void) { return kNoAllocScratchBytes; }

extern "C" int b200cnn_conv2d_bwd_weight_noalloc(const float* dz, const void* x, int x_is_u8, float* dw, float* db,
                                                 int accumulate, const b200cnn_conv_shape* s, b200cnn_stream stream) {
    using namespace b200cnn;
    const size_t need = b200cnn_conv2d_bwd_weight_scratch_bytes(s);  // 0 for an inconsistent shape
    if (need == 0) return set_error(kInvalidValue, "conv2d_bwd_weight_noalloc: inconsistent shape");
    if (need > kNoAllocScratchBytes)
        return set_error(kInvalidValue,
                         "conv2d_bwd_weight_noalloc: this layer needs %zu bytes of scratch but the built-in buffer holds %zu; "
                         "use b200cnn_conv2d_bwd_weight with caller-provided scratch, or a smaller batch",
                         need, kNoAllocScratchBytes);
    float* scratch = nullptr;
    if (int rc = scratch_for_current_device(&scratch)) return rc;
    return launch_conv2d_bwd_weight(dz, x, x_is_u8, dw, db, accumulate, scratch, s, as_stream(stream));
}
