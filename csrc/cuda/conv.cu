// Direct (CUDA-core, fp32) convolution kernels: forward + bias + activation, backward-data fused with
// the previous layer's activation derivative, backward-weight + bias-grad with a deterministic
// two-pass reduction.  These are the generic-shape path of the library (any C/K/stride/pad that fits
// shared memory); the tcgen05 implicit-GEMM path in tc_conv.cu takes over for shapes it supports.
//
// Reference being replaced: Layer_feedForw_conv / Layer_feedBack_conv (/root/reference/cnn.c:175-247)
// and conv_forward_kernel (/root/reference/CUDAcnn.cu:167-195: fp64, one image, no activation).
#include "common.cuh"

namespace b200cnn {

namespace {

constexpr int kConvThreads = 128;
constexpr int kOcTile = 16;  // output channels kept in registers per thread
constexpr int kIcTile = 16;  // input channels kept in registers per thread (backward-data)

template <typename TIn>
__device__ __forceinline__ float load_in(const TIn* p, size_t i, float scale);
template <>
__device__ __forceinline__ float load_in<float>(const float* p, size_t i, float) {
    return __ldg(p + i);
}
template <>
__device__ __forceinline__ float load_in<uint8_t>(const uint8_t* p, size_t i, float scale) {
    return float(__ldg(p + i)) * scale;
}

// One thread = one output pixel (n, oy, ox) x kOcTile output channels.  The weight slice of the
// channel tile sits in shared memory as [c*k*k][kOcTile] so the inner loop reads it as broadcast
// float4s; activations come through the read-only path (every input pixel is reused k*k/stride^2
// times by neighbouring threads, which L1 absorbs).
template <typename TIn>
__global__ void __launch_bounds__(kConvThreads)
conv2d_fwd_kernel(const TIn* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                  float* __restrict__ y, b200cnn_conv_shape s, int act, float in_scale, int nhwc_out) {
    extern __shared__ __align__(16) float ws[];
    const int ckk = s.c * s.k * s.k;
    const int oc0 = blockIdx.y * kOcTile;
    pdl_wait();
    pdl_launch_dependents();
    for (int i = threadIdx.x; i < ckk * kOcTile; i += blockDim.x) {
        const int j = i / kOcTile, o = i % kOcTile;
        ws[i] = (oc0 + o < s.oc) ? __ldg(w + size_t(oc0 + o) * ckk + j) : 0.0f;
    }
    __syncthreads();
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int total = s.n * s.oh * s.ow;
    if (p >= total) return;
    const int ox = p % s.ow, oy = (p / s.ow) % s.oh, n = p / (s.ow * s.oh);

    float acc[kOcTile];
#pragma unroll
    for (int o = 0; o < kOcTile; o++) acc[o] = (oc0 + o < s.oc) ? __ldg(b + oc0 + o) : 0.0f;

    const int y0 = oy * s.stride - s.pad, x0 = ox * s.stride - s.pad;
    for (int c = 0; c < s.c; c++) {
        const size_t plane = (size_t(n) * s.c + c) * s.h * s.w;
        for (int ky = 0; ky < s.k; ky++) {
            const int yy = y0 + ky;
            if (yy < 0 || yy >= s.h) continue;
            for (int kx = 0; kx < s.k; kx++) {
                const int xx = x0 + kx;
                if (xx < 0 || xx >= s.w) continue;
                const float v = load_in<TIn>(x, plane + size_t(yy) * s.w + xx, in_scale);
                const float4* wp = reinterpret_cast<const float4*>(&ws[((c * s.k + ky) * s.k + kx) * kOcTile]);
#pragma unroll
                for (int q = 0; q < kOcTile / 4; q++) {
                    const float4 w4 = wp[q];
                    acc[4 * q + 0] = fmaf(v, w4.x, acc[4 * q + 0]);
                    acc[4 * q + 1] = fmaf(v, w4.y, acc[4 * q + 1]);
                    acc[4 * q + 2] = fmaf(v, w4.z, acc[4 * q + 2]);
                    acc[4 * q + 3] = fmaf(v, w4.w, acc[4 * q + 3]);
                }
            }
        }
    }
    const size_t opix = size_t(oy) * s.ow + ox, oplane = size_t(s.oh) * s.ow;
    if (nhwc_out && oc0 + kOcTile <= s.oc && (s.oc & 3) == 0) {
        // channels innermost: this thread's 16 channels are 64 contiguous bytes
        float4* dst = reinterpret_cast<float4*>(y + (size_t(n) * oplane + opix) * s.oc + oc0);
#pragma unroll
        for (int q = 0; q < kOcTile / 4; q++)
            dst[q] = make_float4(apply_act(act, acc[4 * q]), apply_act(act, acc[4 * q + 1]), apply_act(act, acc[4 * q + 2]),
                                 apply_act(act, acc[4 * q + 3]));
        return;
    }
#pragma unroll
    for (int o = 0; o < kOcTile; o++) {
        if (oc0 + o >= s.oc) continue;
        const size_t idx = nhwc_out ? (size_t(n) * oplane + opix) * s.oc + oc0 + o : (size_t(n) * s.oc + oc0 + o) * oplane + opix;
        y[idx] = apply_act(act, acc[o]);
    }
}

// One thread = one input pixel (n, y, x) x kIcTile input channels; gathers from the <= ceil(k/stride)^2
// output pixels whose window covers it.  Weight slice in shared memory as [oc*k*k][kIcTile].
__global__ void __launch_bounds__(kConvThreads)
conv2d_bwd_data_kernel(const float* __restrict__ dz, const float* __restrict__ w, const float* __restrict__ a_prev,
                       float* __restrict__ dx, b200cnn_conv_shape s, int act_prev) {
    extern __shared__ __align__(16) float ws[];
    const int kk = s.k * s.k;
    const int ic0 = blockIdx.y * kIcTile;
    for (int i = threadIdx.x; i < s.oc * kk * kIcTile; i += blockDim.x) {
        const int c = i % kIcTile, j = i / kIcTile;  // j = oc*kk + tap
        const int oc = j / kk, tap = j % kk;
        ws[i] = (ic0 + c < s.c) ? __ldg(w + (size_t(oc) * s.c + ic0 + c) * kk + tap) : 0.0f;
    }
    __syncthreads();
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int total = s.n * s.h * s.w;
    if (p >= total) return;
    const int xx = p % s.w, yy = (p / s.w) % s.h, n = p / (s.w * s.h);

    float acc[kIcTile];
#pragma unroll
    for (int c = 0; c < kIcTile; c++) acc[c] = 0.0f;

    for (int ky = 0; ky < s.k; ky++) {
        const int ty = yy + s.pad - ky;
        if (ty < 0 || ty % s.stride) continue;
        const int oy = ty / s.stride;
        if (oy >= s.oh) continue;
        for (int kx = 0; kx < s.k; kx++) {
            const int tx = xx + s.pad - kx;
            if (tx < 0 || tx % s.stride) continue;
            const int ox = tx / s.stride;
            if (ox >= s.ow) continue;
            for (int oc = 0; oc < s.oc; oc++) {
                const float g = __ldg(dz + ((size_t(n) * s.oc + oc) * s.oh + oy) * s.ow + ox);
                const float4* wp = reinterpret_cast<const float4*>(&ws[((oc * s.k + ky) * s.k + kx) * kIcTile]);
#pragma unroll
                for (int q = 0; q < kIcTile / 4; q++) {
                    const float4 w4 = wp[q];
                    acc[4 * q + 0] = fmaf(g, w4.x, acc[4 * q + 0]);
                    acc[4 * q + 1] = fmaf(g, w4.y, acc[4 * q + 1]);
                    acc[4 * q + 2] = fmaf(g, w4.z, acc[4 * q + 2]);
                    acc[4 * q + 3] = fmaf(g, w4.w, acc[4 * q + 3]);
                }
            }
        }
    }
#pragma unroll
    for (int c = 0; c < kIcTile; c++) {
        if (ic0 + c >= s.c) continue;
        const size_t idx = ((size_t(n) * s.c + ic0 + c) * s.h + yy) * s.w + xx;
        const float d = a_prev ? act_grad_from_output(act_prev, __ldg(a_prev + idx)) : 1.0f;
        dx[idx] = acc[c] * d;
    }
}

// Backward-weight, pass 1.  A block owns `spb` consecutive samples; per sample it stages the input
// image and the output-gradient image in shared memory, then every thread accumulates its own
// (oc, ic, ky, kx) taps (and the bias taps) in registers.  Result: one partial gradient per block.
constexpr int kBwThreads = 256;

template <typename TIn, int MAXR>
__global__ void __launch_bounds__(kBwThreads)
conv2d_bwd_weight_partial_kernel(const float* __restrict__ dz, const TIn* __restrict__ x,
                                 float* __restrict__ partials, b200cnn_conv_shape s, int spb, float in_scale) {
    extern __shared__ __align__(16) float sm[];
    const int in_sz = s.c * s.h * s.w, out_sz = s.oc * s.oh * s.ow;
    float* xs = sm;
    float* gs = sm + in_sz;
    const int kk = s.k * s.k;
    const int nw = s.oc * s.c * kk, nout = nw + s.oc;

    float acc[MAXR];
#pragma unroll
    for (int r = 0; r < MAXR; r++) acc[r] = 0.0f;

    const int n0 = blockIdx.x * spb;
    const int n1 = min(n0 + spb, s.n);
    for (int n = n0; n < n1; n++) {
        __syncthreads();
        for (int i = threadIdx.x; i < in_sz; i += blockDim.x) xs[i] = load_in<TIn>(x, size_t(n) * in_sz + i, in_scale);
        for (int i = threadIdx.x; i < out_sz; i += blockDim.x) gs[i] = __ldg(dz + size_t(n) * out_sz + i);
        __syncthreads();
#pragma unroll
        for (int r = 0; r < MAXR; r++) {
            const int j = threadIdx.x + r * kBwThreads;
            if (j >= nout) break;
            float a = 0.0f;
            if (j < nw) {
                const int kx = j % s.k, ky = (j / s.k) % s.k, c = (j / kk) % s.c, oc = j / (kk * s.c);
                const float* gp = gs + oc * s.oh * s.ow;
                const float* xp = xs + c * s.h * s.w;
                // Valid output range for this tap (zero padding = skipped terms).
                for (int oy = 0; oy < s.oh; oy++) {
                    const int yy = oy * s.stride - s.pad + ky;
                    if (yy < 0 || yy >= s.h) continue;
                    for (int ox = 0; ox < s.ow; ox++) {
                        const int xx = ox * s.stride - s.pad + kx;
                        if (xx < 0 || xx >= s.w) continue;
                        a = fmaf(gp[oy * s.ow + ox], xp[yy * s.w + xx], a);
                    }
                }
            } else {
                const float* gp = gs + (j - nw) * s.oh * s.ow;
                for (int i = 0; i < s.oh * s.ow; i++) a += gp[i];
            }
            acc[r] += a;
        }
    }
#pragma unroll
    for (int r = 0; r < MAXR; r++) {
        const int j = threadIdx.x + r * kBwThreads;
        if (j < nout) partials[size_t(blockIdx.x) * nout + j] = acc[r];
    }
}

// Backward-weight, pass 2: sum the per-block partials in block order (fixed order => bitwise
// reproducible), write / accumulate into dw and db.
// Block = 32 outputs x 8 chunk-slices: every thread sums its slice of the chunks with all loads in
// flight, then the 8 slice sums are added in slice order through shared memory.  The summation tree is
// fixed, so the result is bitwise reproducible.
constexpr int kRedSlices = 8;
__global__ void __launch_bounds__(32 * kRedSlices)
conv2d_bwd_weight_reduce_kernel(const float* __restrict__ partials, float* __restrict__ dw, float* __restrict__ db,
                                int nchunks, int nw, int nb, int accumulate) {
    __shared__ float red[kRedSlices][32];
    const int lane = threadIdx.x % 32, slice = threadIdx.x / 32;
    const int j = blockIdx.x * 32 + lane;
    const int nout = nw + nb;
    float a = 0.0f;
    pdl_wait();
    pdl_launch_dependents();
    if (j < nout) {
        int c = slice;
        for (; c + 3 * kRedSlices < nchunks; c += 4 * kRedSlices) {
            const float v0 = __ldg(partials + size_t(c) * nout + j);
            const float v1 = __ldg(partials + size_t(c + kRedSlices) * nout + j);
            const float v2 = __ldg(partials + size_t(c + 2 * kRedSlices) * nout + j);
            const float v3 = __ldg(partials + size_t(c + 3 * kRedSlices) * nout + j);
            a += v0;
            a += v1;
            a += v2;
            a += v3;
        }
        for (; c < nchunks; c += kRedSlices) a += __ldg(partials + size_t(c) * nout + j);
    }
    red[slice][lane] = a;
    __syncthreads();
    if (slice != 0 || j >= nout) return;
#pragma unroll
    for (int sl = 1; sl < kRedSlices; sl++) a += red[sl][lane];
    if (j < nw ? dw == nullptr : db == nullptr) return;
    float* dst = j < nw ? dw + j : db + (j - nw);
    *dst = accumulate ? *dst + a : a;
}

int bwd_weight_chunks(const b200cnn_conv_shape* s, int* spb_out) {
    // Aim for ~2 blocks per SM worth of chunks, at least 1 sample per block.
    int spb = ceil_div(s->n, 296);
    if (spb < 1) spb = 1;
    if (spb_out) *spb_out = spb;
    return ceil_div(s->n, spb);
}

bool shape_ok(const b200cnn_conv_shape* s) {
    return s && s->n > 0 && s->c > 0 && s->h > 0 && s->w > 0 && s->oc > 0 && s->oh > 0 && s->ow > 0 && s->k > 0 &&
           s->stride > 0 && s->pad >= 0 && (s->oh - 1) * s->stride + s->k <= s->h + 2 * s->pad &&
           (s->ow - 1) * s->stride + s->k <= s->w + 2 * s->pad;
}

}  // namespace

}  // namespace b200cnn

using namespace b200cnn;

namespace b200cnn {
int reduce_conv_partials(const float* partials, float* dw, float* db, int chunks, int nw, int nb, int accumulate,
                         cudaStream_t st) {
    B200CNN_CUDA(launch_kernel(conv2d_bwd_weight_reduce_kernel, dim3(ceil_div(nw + nb, 32)), dim3(32 * kRedSlices), 0, st, partials, dw, db,
                               chunks, nw, nb, accumulate));
    B200CNN_CHECK_LAUNCH("conv2d_bwd_weight(reduce)");
    return 0;
}
int simt_conv2d_fwd(const void* x, int x_is_u8, const float* w, const float* b, float* y, const b200cnn_conv_shape* s,
                    int act, int nhwc_out, cudaStream_t st);
}  // namespace b200cnn

extern "C" int b200cnn_conv2d_fwd(const void* x, int x_is_u8, const float* w, const float* b, float* y,
                                  const b200cnn_conv_shape* s, int act, b200cnn_stream stream) {
    return simt_conv2d_fwd(x, x_is_u8, w, b, y, s, act, 0, as_stream(stream));
}

int b200cnn::simt_conv2d_fwd(const void* x, int x_is_u8, const float* w, const float* b, float* y,
                             const b200cnn_conv_shape* s, int act, int nhwc_out, cudaStream_t stream) {
    if (!shape_ok(s)) return set_error(kInvalidValue, "conv2d_fwd: inconsistent shape");
    const size_t smem = size_t(s->c) * s->k * s->k * kOcTile * sizeof(float);
    if (smem > 48 * 1024) return set_error(kInvalidValue, "conv2d_fwd: c*k*k = %d too large for the direct kernel", s->c * s->k * s->k);
    const dim3 grid(ceil_div(s->n * s->oh * s->ow, kConvThreads), ceil_div(s->oc, kOcTile));
    if (x_is_u8)
        B200CNN_CUDA(launch_kernel(conv2d_fwd_kernel<uint8_t>, grid, dim3(kConvThreads), smem, stream, static_cast<const uint8_t*>(x),
                                   w, b, y, *s, act, 1.0f / 255.0f, nhwc_out));
    else
        B200CNN_CUDA(launch_kernel(conv2d_fwd_kernel<float>, grid, dim3(kConvThreads), smem, stream, static_cast<const float*>(x), w,
                                   b, y, *s, act, 1.0f, nhwc_out));
    B200CNN_CHECK_LAUNCH("conv2d_fwd");
    return 0;
}

extern "C" int b200cnn_conv2d_bwd_data(const float* dz, const float* w, const float* a_prev, int act_prev,
                                       float* dx, const b200cnn_conv_shape* s, b200cnn_stream stream) {
    if (!shape_ok(s)) return set_error(kInvalidValue, "conv2d_bwd_data: inconsistent shape");
    const size_t smem = size_t(s->oc) * s->k * s->k * kIcTile * sizeof(float);
    if (smem > 48 * 1024) return set_error(kInvalidValue, "conv2d_bwd_data: oc*k*k = %d too large for the direct kernel", s->oc * s->k * s->k);
    const dim3 grid(ceil_div(s->n * s->h * s->w, kConvThreads), ceil_div(s->c, kIcTile));
    conv2d_bwd_data_kernel<<<grid, kConvThreads, smem, as_stream(stream)>>>(dz, w, a_prev, dx, *s, act_prev);
    B200CNN_CHECK_LAUNCH("conv2d_bwd_data");
    return 0;
}

extern "C" size_t b200cnn_conv2d_bwd_weight_scratch_bytes(const b200cnn_conv_shape* s) {
    if (!shape_ok(s)) return 0;
    const size_t nout = size_t(s->oc) * s->c * s->k * s->k + s->oc;
    return size_t(bwd_weight_chunks(s, nullptr)) * nout * sizeof(float);
}

namespace b200cnn {
// Shared by the allocating-caller entry point and any other front end that supplies scratch.
int launch_conv2d_bwd_weight(const float* dz, const void* x, int x_is_u8, float* dw, float* db, int accumulate,
                             float* partials, const b200cnn_conv_shape* s, cudaStream_t st) {
    const int nw = s->oc * s->c * s->k * s->k, nout = nw + s->oc;
    int spb = 1;
    const int chunks = bwd_weight_chunks(s, &spb);
    const size_t smem = (size_t(s->c) * s->h * s->w + size_t(s->oc) * s->oh * s->ow) * sizeof(float);
    if (smem > 200 * 1024) return set_error(kInvalidValue, "conv2d_bwd_weight: one sample (%zu B) does not fit shared memory", smem);
    const int rounds = ceil_div(nout, kBwThreads);
    if (rounds > 32) return set_error(kInvalidValue, "conv2d_bwd_weight: %d weights exceed the direct kernel's limit", nw);

#define LAUNCH_BW(T, R)                                                                                          \
    do {                                                                                                         \
        auto kern = conv2d_bwd_weight_partial_kernel<T, R>;                                                      \
        if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)); \
        kern<<<chunks, kBwThreads, smem, st>>>(dz, static_cast<const T*>(x), partials, *s, spb,                  \
                                               sizeof(T) == 1 ? 1.0f / 255.0f : 1.0f);                           \
    } while (0)
#define DISPATCH_R(T)                         \
    do {                                      \
        if (rounds <= 1) LAUNCH_BW(T, 1);     \
        else if (rounds <= 4) LAUNCH_BW(T, 4); \
        else if (rounds <= 8) LAUNCH_BW(T, 8); \
        else if (rounds <= 20) LAUNCH_BW(T, 20); \
        else LAUNCH_BW(T, 32);                \
    } while (0)
    if (x_is_u8) DISPATCH_R(uint8_t);
    else DISPATCH_R(float);
#undef DISPATCH_R
#undef LAUNCH_BW
    B200CNN_CHECK_LAUNCH("conv2d_bwd_weight(partial)");
    return reduce_conv_partials(partials, dw, db, chunks, nw, s->oc, accumulate, st);
}
}  // namespace b200cnn

extern "C" int b200cnn_conv2d_bwd_weight(const float* dz, const void* x, int x_is_u8, float* dw, float* db,
                                         int accumulate, void* scratch, const b200cnn_conv_shape* s,
                                         b200cnn_stream stream) {
    if (!shape_ok(s)) return set_error(kInvalidValue, "conv2d_bwd_weight: inconsistent shape");
    if (!scratch) return set_error(kInvalidValue, "conv2d_bwd_weight: scratch is NULL");
    return launch_conv2d_bwd_weight(dz, x, x_is_u8, dw, db, accumulate, static_cast<float*>(scratch), s,
                                    as_stream(stream));
}
