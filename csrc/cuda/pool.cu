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

// ---- NHWC variants for the tensor-core convolution chain of the pooled topology -----------------------------------
// x: NHWC [n][h][w][c]; y: NHWC [n][oh][ow][c], or (out_chw) CHW-flattened [n][c*oh*ow] — what a fully connected layer
// reads (/root/reference/cnn.c:116-121).  argmax[n][oh][ow][c] = pixel index (yy * w + xx) of the winner.
namespace b200cnn {
namespace {

__global__ void maxpool_nhwc_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int32_t* __restrict__ argmax, int n,
                                        int c, int h, int w, int oh, int ow, int k, int stride, int out_chw) {
    pdl_wait();
    pdl_launch_dependents();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;  // channel innermost: coalesced loads
    if (i >= n * oh * ow * c) return;
    const int ch = i % c, ox = (i / c) % ow, oy = (i / (c * ow)) % oh, s = i / (c * ow * oh);
    const float* xp = x + size_t(s) * h * w * c + ch;
    int best = -1;
    float bv = 0.0f;
    for (int dy = 0; dy < k; dy++)
        for (int dx = 0; dx < k; dx++) {
            const int idx = (oy * stride + dy) * w + ox * stride + dx;
            const float v = __ldg(xp + size_t(idx) * c);
            if (best < 0 || bv < v) {
                best = idx;
                bv = v;
            }
        }
    const size_t o = out_chw ? (size_t(s) * c + ch) * oh * ow + oy * ow + ox : size_t(i);
    y[o] = bv;
    argmax[i] = best;
}

// dy: NHWC, or CHW-flattened when dy_chw; dx, a_prev: NHWC [n][h][w][c].  Gather form, no atomics.
__global__ void maxpool_nhwc_bwd_kernel(const float* __restrict__ dy, int dy_chw, const int32_t* __restrict__ argmax,
                                        const float* __restrict__ a_prev, int act_prev, float* __restrict__ dx, int n, int c, int h,
                                        int w, int oh, int ow, int k, int stride) {
    pdl_wait();
    pdl_launch_dependents();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * h * w * c) return;
    const int ch = i % c, xx = (i / c) % w, yy = (i / (c * w)) % h, s = i / (c * w * h);
    const int self = yy * w + xx;
    float g = 0.0f;
    const int oy_hi = min(yy / stride, oh - 1), ox_hi = min(xx / stride, ow - 1);
    for (int oy = max(0, (yy - k + stride) / stride); oy <= oy_hi; oy++)
        for (int ox = max(0, (xx - k + stride) / stride); ox <= ox_hi; ox++) {
            const size_t o = ((size_t(s) * oh + oy) * ow + ox) * c + ch;
            if (__ldg(argmax + o) == self) g += __ldg(dy + (dy_chw ? (size_t(s) * c + ch) * oh * ow + oy * ow + ox : o));
        }
    if (a_prev) g *= act_grad_from_output(act_prev, __ldg(a_prev + i));
    dx[i] = g;
}

// One block per sample, for the pool in front of a fully connected layer: the winners are computed in NHWC order
// (coalesced reads, coalesced argmax stores) into shared memory laid out CHW, then the sample's row leaves in one
// coalesced sweep instead of 4-byte stores c*oh*ow floats apart.
__global__ void maxpool_nhwc_fwd_chw_kernel(const float* __restrict__ x, float* __restrict__ y, int32_t* __restrict__ argmax,
                                            int c, int h, int w, int oh, int ow, int k, int stride) {
    extern __shared__ float sy[];  // [c][oh * ow]
    pdl_wait();
    pdl_launch_dependents();
    const int s = blockIdx.x, per = oh * ow * c, opix = oh * ow;
    const float* xs = x + size_t(s) * h * w * c;
    for (int i = threadIdx.x; i < per; i += blockDim.x) {
        const int ch = i % c, ox = (i / c) % ow, oy = i / (c * ow);
        int best = -1;
        float bv = 0.0f;
        if (k == 2) {  // the four loads in flight together (the kernel is a chain of L2 round trips otherwise)
            const int i00 = oy * stride * w + ox * stride;
            const float v0 = __ldg(xs + size_t(i00) * c + ch), v1 = __ldg(xs + size_t(i00 + 1) * c + ch);
            const float v2 = __ldg(xs + size_t(i00 + w) * c + ch), v3 = __ldg(xs + size_t(i00 + w + 1) * c + ch);
            best = i00, bv = v0;
            if (bv < v1) best = i00 + 1, bv = v1;
            if (bv < v2) best = i00 + w, bv = v2;
            if (bv < v3) best = i00 + w + 1, bv = v3;
        } else {
            for (int dy = 0; dy < k; dy++)
                for (int dx = 0; dx < k; dx++) {
                    const int idx = (oy * stride + dy) * w + ox * stride + dx;
                    const float v = __ldg(xs + size_t(idx) * c + ch);
                    if (best < 0 || bv < v) {
                        best = idx;
                        bv = v;
                    }
                }
        }
        sy[ch * opix + oy * ow + ox] = bv;
        argmax[size_t(s) * per + i] = best;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < per; i += blockDim.x) y[size_t(s) * per + i] = sy[i];
}

// Scatter form of the backward pass for non-overlapping windows that tile the input exactly (k == stride, h = oh * k,
// w = ow * k): one thread per POOLED element writes its k * k inputs (the winner gets the gradient, the others zero),
// a quarter of the loads of the gather form.  The activation derivative needs only the winner's activation.
// kChw: dy is CHW-flattened (it comes from a fully connected layer); one block per sample stages the row through
// shared memory so the global reads stay coalesced.
template <bool kChw>
__global__ void maxpool_nhwc_bwd_scatter_kernel(const float* __restrict__ dy, const int32_t* __restrict__ argmax,
                                                const float* __restrict__ a_prev, int act_prev, float* __restrict__ dx, int n,
                                                int c, int h, int w, int oh, int ow, int k) {
    extern __shared__ float sdy[];  // kChw: [c][oh * ow]
    pdl_wait();
    pdl_launch_dependents();
    const int per = oh * ow * c, opix = oh * ow;
    int s, i0, i1, step;
    if (kChw) {
        s = blockIdx.x;
        for (int i = threadIdx.x; i < per; i += blockDim.x) sdy[i] = __ldg(dy + size_t(s) * per + i);
        __syncthreads();
        i0 = threadIdx.x, i1 = per, step = blockDim.x;
    } else {
        const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
        if (gi >= (long long)n * per) return;
        s = int(gi / per);
        i0 = int(gi % per), i1 = i0 + 1, step = 1;
    }
    constexpr int kBatch = 4;  // items per thread whose dependent loads (argmax -> winner's activation) are in flight together
    for (int ib = i0; ib < i1; ib += kBatch * step) {
        int am[kBatch];
        float g[kBatch];
#pragma unroll
        for (int u = 0; u < kBatch; u++) {
            const int i = ib + u * step;
            am[u] = i < i1 ? __ldg(argmax + size_t(s) * per + i) : 0;
        }
#pragma unroll
        for (int u = 0; u < kBatch; u++) {
            const int i = ib + u * step;
            if (i >= i1) continue;
            const int ch = i % c, ox = (i / c) % ow, oy = i / (c * ow);
            g[u] = kChw ? sdy[ch * opix + oy * ow + ox] : __ldg(dy + size_t(s) * per + i);
            if (a_prev) g[u] *= act_grad_from_output(act_prev, __ldg(a_prev + (size_t(s) * h * w + am[u]) * c + ch));
        }
#pragma unroll
        for (int u = 0; u < kBatch; u++) {
            const int i = ib + u * step;
            if (i >= i1) continue;
            const int ch = i % c, ox = (i / c) % ow, oy = i / (c * ow);
            for (int dy_ = 0; dy_ < k; dy_++)
                for (int dx_ = 0; dx_ < k; dx_++) {
                    const int pix = (oy * k + dy_) * w + ox * k + dx_;
                    dx[(size_t(s) * h * w + pix) * c + ch] = pix == am[u] ? g[u] : 0.0f;
                }
        }
    }
}

}  // namespace

int maxpool_nhwc_fwd(const float* x, float* y, int32_t* argmax, int n, int c, int h, int w, int k, int stride, int out_chw,
                     cudaStream_t st) {
    if (n <= 0 || c <= 0 || k <= 0 || stride <= 0 || h < k || w < k || !argmax) return set_error(kInvalidValue, "maxpool_nhwc_fwd: bad shape");
    const int oh = (h - k) / stride + 1, ow = (w - k) / stride + 1;
    const size_t row = size_t(c) * oh * ow * sizeof(float);
    if (out_chw && row <= 48 * 1024) {
        B200CNN_CUDA(launch_kernel(maxpool_nhwc_fwd_chw_kernel, dim3(n), dim3(512), row, st, x, y, argmax, c, h, w, oh, ow, k, stride));
        return 0;
    }
    B200CNN_CUDA(launch_kernel(maxpool_nhwc_fwd_kernel, dim3(ceil_div(n * c * oh * ow, 256)), dim3(256), 0, st, x, y, argmax, n, c, h, w,
                               oh, ow, k, stride, out_chw));
    return 0;
}

int maxpool_nhwc_bwd(const float* dy, int dy_chw, const int32_t* argmax, const float* a_prev, int act_prev, float* dx, int n, int c,
                     int h, int w, int k, int stride, cudaStream_t st) {
    if (n <= 0 || c <= 0 || k <= 0 || stride <= 0 || h < k || w < k || !argmax) return set_error(kInvalidValue, "maxpool_nhwc_bwd: bad shape");
    const int oh = (h - k) / stride + 1, ow = (w - k) / stride + 1;
    const size_t row = size_t(c) * oh * ow * sizeof(float);
    if (k == stride && h == oh * k && w == ow * k && (!dy_chw || row <= 48 * 1024)) {
        if (dy_chw)
            B200CNN_CUDA(launch_kernel(maxpool_nhwc_bwd_scatter_kernel<true>, dim3(n), dim3(512), row, st, dy, argmax, a_prev, act_prev, dx,
                                       n, c, h, w, oh, ow, k));
        else
            B200CNN_CUDA(launch_kernel(maxpool_nhwc_bwd_scatter_kernel<false>, dim3(ceil_div(n * c * oh * ow, 256)), dim3(256), 0, st, dy,
                                       argmax, a_prev, act_prev, dx, n, c, h, w, oh, ow, k));
        return 0;
    }
    B200CNN_CUDA(launch_kernel(maxpool_nhwc_bwd_kernel, dim3(ceil_div(n * c * h * w, 256)), dim3(256), 0, st, dy, dy_chw, argmax, a_prev,
                               act_prev, dx, n, c, h, w, oh, ow, k, stride));
    return 0;
}

}  // namespace b200cnn

extern "C" int b200cnn_maxpool_nhwc_fwd(const float* x, float* y, int32_t* argmax, int n, int c, int h, int w, int k, int stride,
                                        int out_chw, b200cnn_stream stream) {
    return b200cnn::maxpool_nhwc_fwd(x, y, argmax, n, c, h, w, k, stride, out_chw, b200cnn::as_stream(stream));
}
extern "C" int b200cnn_maxpool_nhwc_bwd(const float* dy, int dy_chw, const int32_t* argmax, const float* a_prev, int act_prev, float* dx,
                                        int n, int c, int h, int w, int k, int stride, b200cnn_stream stream) {
    return b200cnn::maxpool_nhwc_bwd(dy, dy_chw, argmax, a_prev, act_prev, dx, n, c, h, w, k, stride, b200cnn::as_stream(stream));
}
