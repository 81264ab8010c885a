// Plain SGD (single GPU).  The multi-GPU form of the same maths, fused with the gradient
// reduction over NVLink, is csrc/comm/allreduce_sgd.cu.
// Reference being replaced: Layer_update (/root/reference/cnn.c:303-314), a CPU loop.
#include "common.cuh"

namespace b200cnn {
namespace {

__global__ void sgd_kernel(float* __restrict__ w, float* __restrict__ g, size_t n, float lr, int zero_grad) {
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    const size_t n4 = n / 4;
    float4* w4 = reinterpret_cast<float4*>(w);
    float4* g4 = reinterpret_cast<float4*>(g);
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 a = w4[i];
        const float4 d = g4[i];
        a.x -= lr * d.x;
        a.y -= lr * d.y;
        a.z -= lr * d.z;
        a.w -= lr * d.w;
        w4[i] = a;
        if (zero_grad) g4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (size_t i = n4 * 4 + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        w[i] -= lr * g[i];
        if (zero_grad) g[i] = 0.0f;
    }
}

}  // namespace
}  // namespace b200cnn

using namespace b200cnn;

extern "C" int b200cnn_sgd_update(float* w, float* g, size_t n, float lr, int zero_grad, b200cnn_stream stream) {
    if (n == 0) return 0;
    if ((reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(g)) & 15)
        return set_error(kInvalidValue, "sgd_update: buffers must be 16-byte aligned");
    const int blocks = int(std::min<size_t>((n / 4 + 255) / 256 + 1, 148 * 4));
    sgd_kernel<<<blocks, 256, 0, as_stream(stream)>>>(w, g, n, lr, zero_grad);
    B200CNN_CHECK_LAUNCH("sgd_update");
    return 0;
}
