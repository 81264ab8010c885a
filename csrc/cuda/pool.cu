// Max pooling forward / backward (K7).  The reference's network has no pooling layer (it
// down-samples with stride-2 convolutions, /root/reference/cnn.c:419-422); the project's stated
// topology does, so the layer type exists on both the CPU and GPU engines.
#include "common.cuh"

namespace b200cnn {
namespace {

__global__ void maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int32_t* __restrict__ argmax,
                                   int nc, int h, int w, int oh, int ow, int k, int stride) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nc * oh * ow) return;
    const int ox = i % ow, oy = (i / ow) % oh, p = i / (ow * oh);
    const float* xp = x + size_t(p) * h * w;
    int best = -1;
    float bv = 0.0f;
    for (int dy = 0; dy < k; dy++)
        for (int dx = 0; dx < k; dx++) {
            const int idx = (oy * stride + dy) * w + ox * stride + dx;
            const float v = __ldg(xp + idx);
            if (best < 0 || bv < v) {
                best = idx;
                bv = v;
            }
        }
    y[i] = bv;
    if (argmax) argmax[i] = best;
}

// Gather form (no atomics): each input pixel looks at the <= ceil(k/stride)^2 windows covering it.
__global__ void maxpool_bwd_kernel(const float* __restrict__ dy, const int32_t* __restrict__ argmax,
                                   const float* __restrict__ a_prev, int act_prev, float* __restrict__ dx, int nc,
                                   int h, int w, int oh, int ow, int k, int stride) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nc * h * w) return;
    const int xx = i % w, yy = (i / w) % h, p = i / (w * h);
    const int self = yy * w + xx;
    float g = 0.0f;
    const int oy_hi = min(yy / stride, oh - 1), ox_hi = min(xx / stride, ow - 1);
    for (int oy = max(0, (yy - k + stride) / stride); oy <= oy_hi; oy++)
        for (int ox = max(0, (xx - k + stride) / stride); ox <= ox_hi; ox++) {
            const size_t o = (size_t(p) * oh + oy) * ow + ox;
            if (__ldg(argmax + o) == self) g += __ldg(dy + o);
        }
    if (a_prev) g *= act_grad_from_output(act_prev, __ldg(a_prev + i));
    dx[i] = g;
}

}  // namespace
}  // namespace b200cnn

using namespace b200cnn;

extern "C" int b200cnn_maxpool_fwd(const float* x, float* y, int32_t* argmax, int n, int c, int h, int w, int k,
                                   int stride, b200cnn_stream stream) {
    if (n <= 0 || c <= 0 || k <= 0 || stride <= 0 || h < k || w < k) return set_error(kInvalidValue, "maxpool_fwd: bad shape");
    const int oh = (h - k) / stride + 1, ow = (w - k) / stride + 1;
    const int total = n * c * oh * ow;
    maxpool_fwd_kernel<<<ceil_div(total, 256), 256, 0, as_stream(stream)>>>(x, y, argmax, n * c, h, w, oh, ow, k, stride);
    B200CNN_CHECK_LAUNCH("maxpool_fwd");
    return 0;
}

extern "C" int b200cnn_maxpool_bwd(const float* dy, const int32_t* argmax, const float* a_prev, int act_prev,
                                   float* dx, int n, int c, int h, int w, int k, int stride, b200cnn_stream stream) {
    if (n <= 0 || c <= 0 || k <= 0 || stride <= 0 || h < k || w < k || !argmax) return set_error(kInvalidValue, "maxpool_bwd: bad shape");
    const int oh = (h - k) / stride + 1, ow = (w - k) / stride + 1;
    const int total = n * c * h * w;
    maxpool_bwd_kernel<<<ceil_div(total, 256), 256, 0, as_stream(stream)>>>(dy, argmax, a_prev, act_prev, dx, n * c, h, w, oh, ow, k, stride);
    B200CNN_CHECK_LAUNCH("maxpool_bwd");
    return 0;
}
