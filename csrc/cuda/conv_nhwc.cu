// CUDA-core kernels of the NHWC convolution chain that do not map onto tcgen05 with TF32 operands:
//   * backward-weight: the GEMM is dW[oc][c,tap] = sum_pixels dz[pixel][oc] * patch[pixel][c,tap], i.e. both
//     operands are "pixel-major".  TF32 M/N-major UMMA operands exist only in the 128B/32B-atom layout,
//     which needs 32 contiguous floats per pixel per operand; a 16-channel tap slice has 16.  So these
//     stay register-tiled FMA kernels (4 oc x 4 c per thread, float4 shared-memory reads), ~58 MFLOP.
//   * the first convolution's backward-weight (1 input channel, 144 weights).
// Both write per-block partial sums that conv2d_bwd_weight_reduce sums in a fixed order.
// Replaces the weight half of Layer_feedBack_conv (/root/reference/cnn.c:236, :242).
#include <algorithm>

#include <cstdlib>

#include "backend.cuh"
#include "common.cuh"

namespace b200cnn {
namespace {

template <typename TIn>
__device__ __forceinline__ float cvt_in(TIn v, float scale);
template <>
__device__ __forceinline__ float cvt_in<float>(float v, float) { return v; }
template <>
__device__ __forceinline__ float cvt_in<uint8_t>(uint8_t v, float scale) { return float(v) * scale; }

// Valid output range [lo, hi) along one axis for filter offset k: 0 <= o*stride - pad + k < in.
__device__ __forceinline__ void tap_range(int k, int stride, int pad, int in, int out, int& lo, int& hi) {
    lo = max(0, (pad - k + stride - 1) / stride);
    hi = min(out, (in - 1 + pad - k) / stride + 1);
}

// x NHWC [n][h][w][C], dz NHWC or CHW [n][..] (dz_chw); thread = (4 oc, 4 c, 1 tap).  All `spb` samples of the
// block are staged in shared memory in ONE load phase (the kernel is latency-, not throughput-bound).
__global__ void __launch_bounds__(512)
conv_bwd_weight_nhwc_kernel(const float* __restrict__ x, const float* __restrict__ dz, float* __restrict__ partials,
                            b200cnn_conv_shape s, int spb, int dz_chw) {
    extern __shared__ __align__(16) float sm[];
    const int C = s.c, OC = s.oc, kk = s.k * s.k;
    const int pix = s.oh * s.ow;
    const int in_sz = s.h * s.w * C, out_sz = pix * OC;
    const int c4n = C / 4, oc4n = OC / 4;
    const int t = threadIdx.x;
    const int nthr_work = oc4n * c4n * kk;
    const bool worker = t < nthr_work;
    const int oc4 = t % oc4n, c4 = (t / oc4n) % c4n, tap = t / (oc4n * c4n);
    const int ky = tap / s.k, kx = tap % s.k;
    const bool bias_thread = worker && c4 == 0;  // the kk tap-threads of each oc4 share the bias sum (pixel p = tap mod kk)

    const int n0 = blockIdx.x * spb, ns = min(spb, s.n - n0);
    pdl_wait();
    pdl_launch_dependents();
    float* xs = sm;                       // [ns][h][w][C]
    float* gs = sm + size_t(spb) * in_sz; // [ns][pix][OC]
    {
        const float4* xg = reinterpret_cast<const float4*>(x + size_t(n0) * in_sz);
        for (int i = t; i < ns * in_sz / 4; i += blockDim.x) reinterpret_cast<float4*>(xs)[i] = __ldg(xg + i);
        if (!dz_chw) {
            const float4* gg = reinterpret_cast<const float4*>(dz + size_t(n0) * out_sz);
            for (int i = t; i < ns * out_sz / 4; i += blockDim.x) reinterpret_cast<float4*>(gs)[i] = __ldg(gg + i);
        } else {
            // CHW -> HWC while staging: coalesced reads along the pixel index; (p, oc, nn) advance incrementally
            // instead of three runtime divisions per element.
            int p = t % pix, oc = (t / pix) % OC, nn = t / out_sz;
            const int dp = int(blockDim.x) % pix, doc = int(blockDim.x) / pix;
            for (int i = t; i < ns * out_sz; i += blockDim.x) {
                gs[(nn * pix + p) * OC + oc] = __ldg(dz + size_t(n0) * out_sz + i);
                p += dp;
                oc += doc;
                if (p >= pix) { p -= pix; oc++; }
                while (oc >= OC) { oc -= OC; nn++; }
            }
        }
    }
    __syncthreads();

    float acc[4][4];
    float bacc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
    int oy0, oy1, ox0, ox1;
    tap_range(ky, s.stride, s.pad, s.h, s.oh, oy0, oy1);
    tap_range(kx, s.stride, s.pad, s.w, s.ow, ox0, ox1);

    for (int nn = 0; worker && nn < ns; nn++) {
        const float* xsn = xs + nn * in_sz;
        const float* gsn = gs + nn * out_sz;
        if (bias_thread)
            for (int p = tap; p < pix; p += kk) {
                const float4 g = *reinterpret_cast<const float4*>(gsn + p * OC + oc4 * 4);
                bacc[0] += g.x; bacc[1] += g.y; bacc[2] += g.z; bacc[3] += g.w;
            }
        for (int oy = oy0; oy < oy1; oy++) {
            const float* xrow = xsn + ((oy * s.stride - s.pad + ky) * s.w - s.pad + kx) * C + c4 * 4;
            const float* grow = gsn + oy * s.ow * OC + oc4 * 4;
#pragma unroll 4
            for (int ox = ox0; ox < ox1; ox++) {
                const float4 g = *reinterpret_cast<const float4*>(grow + ox * OC);
                const float4 a = *reinterpret_cast<const float4*>(xrow + ox * s.stride * C);
                acc[0][0] = fmaf(g.x, a.x, acc[0][0]); acc[0][1] = fmaf(g.x, a.y, acc[0][1]);
                acc[0][2] = fmaf(g.x, a.z, acc[0][2]); acc[0][3] = fmaf(g.x, a.w, acc[0][3]);
                acc[1][0] = fmaf(g.y, a.x, acc[1][0]); acc[1][1] = fmaf(g.y, a.y, acc[1][1]);
                acc[1][2] = fmaf(g.y, a.z, acc[1][2]); acc[1][3] = fmaf(g.y, a.w, acc[1][3]);
                acc[2][0] = fmaf(g.z, a.x, acc[2][0]); acc[2][1] = fmaf(g.z, a.y, acc[2][1]);
                acc[2][2] = fmaf(g.z, a.z, acc[2][2]); acc[2][3] = fmaf(g.z, a.w, acc[2][3]);
                acc[3][0] = fmaf(g.w, a.x, acc[3][0]); acc[3][1] = fmaf(g.w, a.y, acc[3][1]);
                acc[3][2] = fmaf(g.w, a.z, acc[3][2]); acc[3][3] = fmaf(g.w, a.w, acc[3][3]);
            }
        }
    }
    const int nw = OC * C * kk, nout = nw + OC;
    float* out = partials + size_t(blockIdx.x) * nout;
    if (worker) {
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) out[((oc4 * 4 + i) * C + c4 * 4 + j) * kk + tap] = acc[i][j];  // OIHW
    }
    // bias: the kk tap-threads of an oc4 group add their pixel slices in tap order through shared memory
    __syncthreads();  // every worker is past its reads of xs / gs: reuse the front of xs
    float* bred = xs;  // [kk][OC]
    if (bias_thread)
#pragma unroll
        for (int i = 0; i < 4; i++) bred[tap * OC + oc4 * 4 + i] = bacc[i];
    __syncthreads();
    if (t < OC) {
        float a = 0.f;
        for (int k2 = 0; k2 < kk; k2++) a += bred[k2 * OC + t];
        out[nw + t] = a;
    }
}

// Single input channel: x [n][h][w] (uint8 or float), dz NHWC [n][oh][ow][OC].
// Work item = (4 output channels, 1 tap) or (4 output channels of the bias); kC1Threads / items pixel-row slices
// work on each item and are summed through shared memory.  Inner loop: one float4 of dz + one pixel -> 4 FMAs.
constexpr int kC1Threads = 640;

// Instrumentation (tools/diag_c1.py; off unless b200cnn_debug_c1_enable(1)): clock64 stamps of block 0 [0..7] and of
// the block that ran the final reduction [8..15].
__device__ long long g_c1_stamps[16];
__device__ int g_c1_debug;
#define C1_STAMP(slot) do { if (g_c1_debug && t == 0) g_c1_stamps[(slot)] = clock64(); } while (0)

// Tail shared by the single-channel weight-gradient kernels: every block has stored its `nout` partial sums.
// `red`: shared memory, (blockDim.x / nout) * nout floats.
__device__ __forceinline__ void last_block_reduce(const float* __restrict__ partials, int nout, int nw, float* red,
                                                  unsigned* __restrict__ counter, float* __restrict__ dw, float* __restrict__ db,
                                                  int accumulate, float* __restrict__ sgd_w, float* __restrict__ sgd_b,
                                                  const float* __restrict__ lr_dev) {
    const int t = threadIdx.x;
    // The block that finishes last sums all partials in block order (fixed order => reproducible) and
    // writes the gradient: no second kernel on the tail of the backward pass.
    // Release / acquire around the counter by ONE thread (cumulative over the CTA barrier) instead of a sequentially
    // consistent __threadfence() in every thread: MEMBAR.SC.GPU x 640 threads was ~1 us of this kernel's tail.
    __shared__ unsigned is_last;
    __syncthreads();
    if (blockIdx.x == 0) C1_STAMP(3);
    if (t == 0) {
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
        const unsigned done = atomicInc(counter, gridDim.x - 1);  // wraps back to 0 for the next call
        is_last = (done == gridDim.x - 1);
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
    __syncthreads();
    if (blockIdx.x == 0) C1_STAMP(4);
    if (!is_last) return;
    C1_STAMP(8);
    const int j = t % nout, part = t / nout, parts = int(blockDim.x) / nout;
    if (part < parts) {
        // All of a thread's L2 loads in flight at once where possible (32, then 8, then 1): every batch is one
        // ~0.8 us round trip on the step's critical path, whatever it carries.
        float a = 0.f;
        int c = part;
        for (; c + 31 * parts < int(gridDim.x); c += 32 * parts) {
            float v[32];
#pragma unroll
            for (int u = 0; u < 32; u++) v[u] = __ldcg(partials + size_t(c + u * parts) * nout + j);
#pragma unroll
            for (int u = 0; u < 32; u++) a += v[u];
        }
        for (; c + 7 * parts < int(gridDim.x); c += 8 * parts) {
            float v[8];
#pragma unroll
            for (int u = 0; u < 8; u++) v[u] = __ldcg(partials + size_t(c + u * parts) * nout + j);
#pragma unroll
            for (int u = 0; u < 8; u++) a += v[u];
        }
        for (; c < int(gridDim.x); c += parts) a += __ldcg(partials + size_t(c) * nout + j);
        red[part * nout + j] = a;
    }
    __syncthreads();
    C1_STAMP(9);
    if (t < nout) {
        float a = red[t];
        for (int sl = 1; sl < parts; sl++) a += red[sl * nout + t];
        float* dst = t < nw ? (dw ? dw + t : nullptr) : (db ? db + (t - nw) : nullptr);
        if (dst) {
            a = accumulate ? *dst + a : a;
            *dst = a;
        }
        // Optional fused SGD for this layer: being the last kernel of backward, applying the update here takes a
        // whole kernel off the end of the step's critical path.
        if (sgd_w) {
            float* wdst = t < nw ? sgd_w + t : sgd_b + (t - nw);
            *wdst -= *lr_dev * a;
        }
    }
    C1_STAMP(10);
}

template <typename TIn>
__global__ void __launch_bounds__(kC1Threads)
conv_bwd_weight_c1_nhwc_kernel(const TIn* __restrict__ x, const float* __restrict__ dz, float* __restrict__ partials,
                               b200cnn_conv_shape s, int spb, float in_scale, float* __restrict__ dw, float* __restrict__ db,
                               int accumulate, unsigned* __restrict__ counter, float* __restrict__ sgd_w,
                               float* __restrict__ sgd_b, const float* __restrict__ lr_dev) {
    extern __shared__ __align__(16) float sm[];
    const int OC = s.oc, kk = s.k * s.k, pix = s.oh * s.ow, oc4n = OC / 4;
    const int in_sz = s.h * s.w, in_pad = (in_sz + 3) & ~3, out_sz = pix * OC;
    const int t = threadIdx.x;
    const int nw = OC * kk, nout = nw + OC;
    const int items = oc4n * (kk + 1), slices = kC1Threads / items;
    const int n0 = blockIdx.x * spb, ns = min(spb, s.n - n0);
    float* xs = sm;                         // [ns][in_pad]
    float* gs = sm + size_t(spb) * in_pad;  // [ns][pix][OC]
    float* red = gs + size_t(spb) * out_sz; // [slices][nout]
    if (blockIdx.x == 0) C1_STAMP(0);
    pdl_wait();
    pdl_launch_dependents();
    if (blockIdx.x == 0) C1_STAMP(1);
    if (sizeof(TIn) == 1 && (in_sz & 3) == 0) {
        // four pixels per load
        const uchar4* x4 = reinterpret_cast<const uchar4*>(x + size_t(n0) * in_sz);
        for (int i = t; i < ns * in_sz / 4; i += kC1Threads) {
            const uchar4 v = __ldg(x4 + i);
            const int nn = (4 * i) / in_sz, off = (4 * i) % in_sz;
            float* d = xs + nn * in_pad + off;
            d[0] = v.x * in_scale; d[1] = v.y * in_scale; d[2] = v.z * in_scale; d[3] = v.w * in_scale;
        }
    } else {
        for (int i = t; i < ns * in_sz; i += kC1Threads)
            xs[(i / in_sz) * in_pad + i % in_sz] = cvt_in<TIn>(__ldg(x + size_t(n0) * in_sz + i), in_scale);
    }
    {
        const float4* gg = reinterpret_cast<const float4*>(dz + size_t(n0) * out_sz);
        for (int i = t; i < ns * out_sz / 4; i += kC1Threads) reinterpret_cast<float4*>(gs)[i] = __ldg(gg + i);
    }
    __syncthreads();
    if (blockIdx.x == 0) C1_STAMP(2);
    const int item = t % items, slice = t / items;
    if (slice < slices) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        const int oc4 = item % oc4n, tap = item / oc4n;  // tap == kk: the bias item
        if (tap < kk) {
            const int ky = tap / s.k, kx = tap % s.k;
            int oy0, oy1, ox0, ox1;
            tap_range(ky, s.stride, s.pad, s.h, s.oh, oy0, oy1);
            tap_range(kx, s.stride, s.pad, s.w, s.ow, ox0, ox1);
            const int nrows = oy1 - oy0;
            for (int rr = slice; rr < ns * nrows; rr += slices) {
                const int nn = rr / nrows, oy = oy0 + rr % nrows;
                const float* xp = xs + nn * in_pad + (oy * s.stride - s.pad + ky) * s.w - s.pad + kx + ox0 * s.stride;
                const float4* gp = reinterpret_cast<const float4*>(gs + (nn * pix + oy * s.ow + ox0) * OC) + oc4;
                for (int ox = ox0; ox < ox1; ox++, xp += s.stride, gp += oc4n) {
                    const float4 g = *gp;
                    const float xv = *xp;
                    acc[0] = fmaf(g.x, xv, acc[0]);
                    acc[1] = fmaf(g.y, xv, acc[1]);
                    acc[2] = fmaf(g.z, xv, acc[2]);
                    acc[3] = fmaf(g.w, xv, acc[3]);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; i++) red[slice * nout + (oc4 * 4 + i) * kk + tap] = acc[i];  // OIHW, 1 input channel
        } else {
            for (int p = slice; p < ns * pix; p += slices) {
                const float4 g = *(reinterpret_cast<const float4*>(gs + p * OC) + oc4);
                acc[0] += g.x; acc[1] += g.y; acc[2] += g.z; acc[3] += g.w;
            }
#pragma unroll
            for (int i = 0; i < 4; i++) red[slice * nout + nw + oc4 * 4 + i] = acc[i];
        }
    }
    __syncthreads();
    if (t < nout) {
        float a = 0.f;
        for (int sl = 0; sl < slices; sl++) a += red[sl * nout + t];
        partials[size_t(blockIdx.x) * nout + t] = a;
    }
    last_block_reduce(partials, nout, nw, red, counter, dw, db, accumulate, sgd_w, sgd_b, lr_dev);
}

// ---- first convolution of the pooled topology: conv (1 input channel, 3x3, stride 1, pad 1) + activation + 2x2 max pool ----
// Forward as ONE kernel: the un-pooled activations (4x the pooled ones, 12.8 MB per 256 MNIST samples) never have to
// leave the SM.  One block per sample; a work item = one pooled pixel x 4 output channels: a 4x4 input patch in
// registers, the 2x2 window's four convolutions, then the max in the scan order of the separate pool kernel (first
// maximum wins).  a_full (optional): the un-pooled activations, NHWC, for callers that want to look at them.
// Replaces Layer_feedForw_conv (/root/reference/cnn.c:175-211) for this layer plus the pool kernel behind it.
constexpr int kCpThreads = 256;

template <typename TIn>
__global__ void __launch_bounds__(kCpThreads)
conv1_pool_fwd_kernel(const TIn* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                      float* __restrict__ a_full, float* __restrict__ yp, int32_t* __restrict__ argmax, b200cnn_conv_shape s,
                      int act, float in_scale) {
    extern __shared__ __align__(16) float sm[];
    const int OC = s.oc, oc4n = OC / 4, wp = s.w + 2, ph = s.oh / 2, pw = s.ow / 2;
    float* xs = sm;                               // [(h + 2)][(w + 2)], zero border
    float* ws = sm + (((s.h + 2) * wp + 3) & ~3);  // [9][OC]
    float* bs = ws + 9 * OC;                      // [OC]
    const int t = threadIdx.x, n = blockIdx.x;
    for (int i = t; i < (s.h + 2) * wp; i += kCpThreads) {
        const int yy = i / wp - 1, xx = i % wp - 1;
        if (yy < 0 || yy >= s.h || xx < 0 || xx >= s.w) xs[i] = 0.0f;
    }
    pdl_wait();  // (the weights too may still be under the previous step's update)
    pdl_launch_dependents();
    for (int i = t; i < 9 * OC; i += kCpThreads) ws[i] = __ldg(w + (i % OC) * 9 + i / OC);  // OIHW -> [tap][oc]
    for (int i = t; i < OC; i += kCpThreads) bs[i] = __ldg(b + i);
    for (int i = t; i < s.h * s.w; i += kCpThreads)
        xs[(i / s.w + 1) * wp + i % s.w + 1] = cvt_in<TIn>(__ldg(x + size_t(n) * s.h * s.w + i), in_scale);
    __syncthreads();
    for (int item = t; item < ph * pw * oc4n; item += kCpThreads) {
        const int oc4 = item % oc4n, q = item / oc4n, py = q / pw, px = q % pw;
        float v[4][4];
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int c = 0; c < 4; c++) v[r][c] = xs[(2 * py + r) * wp + 2 * px + c];
        const float4 b4 = *reinterpret_cast<const float4*>(bs + oc4 * 4);
        float best[4] = {0.f, 0.f, 0.f, 0.f};
        int arg[4] = {-1, -1, -1, -1};
#pragma unroll
        for (int dy = 0; dy < 2; dy++)
#pragma unroll
            for (int dx = 0; dx < 2; dx++) {
                float acc[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
                for (int ky = 0; ky < 3; ky++)
#pragma unroll
                    for (int kx = 0; kx < 3; kx++) {
                        const float4 w4 = *reinterpret_cast<const float4*>(ws + (ky * 3 + kx) * OC + oc4 * 4);
                        const float xv = v[dy + ky][dx + kx];
                        acc[0] = fmaf(xv, w4.x, acc[0]);
                        acc[1] = fmaf(xv, w4.y, acc[1]);
                        acc[2] = fmaf(xv, w4.z, acc[2]);
                        acc[3] = fmaf(xv, w4.w, acc[3]);
                    }
                const int pix = (2 * py + dy) * s.ow + 2 * px + dx;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    acc[j] = apply_act(act, acc[j]);
                    if (arg[j] < 0 || best[j] < acc[j]) {
                        best[j] = acc[j];
                        arg[j] = pix;
                    }
                }
                if (a_full)
                    *reinterpret_cast<float4*>(a_full + (size_t(n) * s.oh * s.ow + pix) * OC + oc4 * 4) =
                        make_float4(acc[0], acc[1], acc[2], acc[3]);
            }
        const size_t o = (size_t(n) * ph * pw + q) * OC + oc4 * 4;
        *reinterpret_cast<float4*>(yp + o) = make_float4(best[0], best[1], best[2], best[3]);
        *reinterpret_cast<int4*>(argmax + o) = make_int4(arg[0], arg[1], arg[2], arg[3]);
    }
}

// Weight gradient of the same layer straight from the POOLED gradient: only the winner of every 2x2 window carries a
// gradient, so
//     dW[oc][ky][kx] = sum over (n, pooled pixel q) of g * x[n][wy - 1 + ky][wx - 1 + kx],   (wy, wx) = argmax(n, q, oc),
//     g = dyp[n][q][oc] * act'(yp[n][q][oc])        (the winner's activation IS the pooled output)
// — a quarter of the products of the dense form, and the un-pooled gradient (what a pool-backward kernel would
// write: 12.8 MB per 256 samples) is never materialised.  Thread = (output channel, every (blockDim / OC)-th pooled
// pixel): 9 + 1 register accumulators, reduced over the block through shared memory, then over the blocks by the last
// one to finish (last_block_reduce).  Replaces the weight half of Layer_feedBack_conv (/root/reference/cnn.c:236, :242).
constexpr int kCpwThreads = 512;

template <typename TIn>
__global__ void __launch_bounds__(kCpwThreads)
conv1_pool_bwd_weight_kernel(const TIn* __restrict__ x, const float* __restrict__ dyp, const float* __restrict__ yp,
                             const int32_t* __restrict__ argmax, int act, float* __restrict__ partials, b200cnn_conv_shape s,
                             int spb, float in_scale, float* __restrict__ dw, float* __restrict__ db, int accumulate,
                             unsigned* __restrict__ counter, float* __restrict__ sgd_w, float* __restrict__ sgd_b,
                             const float* __restrict__ lr_dev) {
    extern __shared__ __align__(16) float sm[];
    const int OC = s.oc, wp = s.w + 2, img = ((s.h + 2) * wp + 3) & ~3, per = (s.oh / 2) * (s.ow / 2) * OC;
    const int nw = OC * 9, nout = nw + OC, groups = kCpwThreads / OC;
    const int t = threadIdx.x, n0 = blockIdx.x * spb, ns = min(spb, s.n - n0);
    float* xs = sm;                      // [spb][(h + 2)][(w + 2)], zero border
    float* red = sm + size_t(spb) * img;  // [groups][nout]
    for (int i = t; i < spb * img; i += kCpwThreads) xs[i] = 0.0f;
    __syncthreads();
    pdl_wait();
    pdl_launch_dependents();
    for (int i = t; i < ns * s.h * s.w; i += kCpwThreads) {
        const int nn = i / (s.h * s.w), r = i % (s.h * s.w);
        xs[nn * img + (r / s.w + 1) * wp + r % s.w + 1] = cvt_in<TIn>(__ldg(x + size_t(n0) * s.h * s.w + i), in_scale);
    }
    __syncthreads();
    float acc[9], bacc = 0.f;
#pragma unroll
    for (int k2 = 0; k2 < 9; k2++) acc[k2] = 0.f;
    const size_t base = size_t(n0) * per;
    for (int i = t; i < ns * per; i += kCpwThreads) {  // kCpwThreads % OC == 0: a thread keeps its output channel
        float g = __ldg(dyp + base + i);
        g *= act_grad_from_output(act, __ldg(yp + base + i));
        const int am = __ldg(argmax + base + i);
        const float* xp = xs + (i / per) * img + (am / s.ow) * wp + am % s.ow;  // top-left tap in padded coordinates
        bacc += g;
#pragma unroll
        for (int ky = 0; ky < 3; ky++)
#pragma unroll
            for (int kx = 0; kx < 3; kx++) acc[ky * 3 + kx] = fmaf(g, xp[ky * wp + kx], acc[ky * 3 + kx]);
    }
    const int oc = t % OC, grp = t / OC;
#pragma unroll
    for (int k2 = 0; k2 < 9; k2++) red[grp * nout + oc * 9 + k2] = acc[k2];  // OIHW, 1 input channel
    red[grp * nout + nw + oc] = bacc;
    __syncthreads();
    if (t < nout) {
        float a = 0.f;
        for (int g2 = 0; g2 < groups; g2++) a += red[g2 * nout + t];
        partials[size_t(blockIdx.x) * nout + t] = a;
    }
    last_block_reduce(partials, nout, nw, red, counter, dw, db, accumulate, sgd_w, sgd_b, lr_dev);
}

}  // namespace

extern "C" int b200cnn_debug_c1_enable(int on) {
    return cudaMemcpyToSymbol(g_c1_debug, &on, sizeof(int)) == cudaSuccess ? 0 : 1;
}
extern "C" int b200cnn_debug_c1_stamps(long long* out16) {
    return cudaMemcpyFromSymbol(out16, g_c1_stamps, 16 * sizeof(long long)) == cudaSuccess ? 0 : 1;
}

int reduce_conv_partials(const float* partials, float* dw, float* db, int chunks, int nw, int nb, int accumulate,
                         cudaStream_t st);

bool nhwc_bwd_weight_supported(const b200cnn_conv_shape* s, int x_is_u8) {
    if (s->c == 1) return s->oc * s->k * s->k + s->oc <= kC1Threads && s->oc % 4 == 0;
    return !x_is_u8 && s->c % 4 == 0 && s->oc % 4 == 0 && (s->oc / 4) * (s->c / 4) * s->k * s->k <= 512;
}

int nhwc_bwd_weight_chunks(int n, int* spb_out);
int nhwc_bwd_weight_chunks_for(const b200cnn_conv_shape* s, int* spb_out);

// Dynamic shared memory of the CUDA-core weight-gradient kernels for a batch of s->n samples (every sample of a block is
// staged at once); a shape is usable only while this stays under the 200 KB the launcher allows.
size_t nhwc_bwd_weight_smem_bytes(const b200cnn_conv_shape* s) {
    int spb = 1;
    nhwc_bwd_weight_chunks_for(s, &spb);
    const size_t in_floats = s->c == 1 ? size_t((s->h * s->w + 3) & ~3) : size_t(s->h) * s->w * s->c;
    return size_t(spb) * (in_floats + size_t(s->oh) * s->ow * s->oc) * sizeof(float) + (s->c == 1 ? 2 * kC1Threads * sizeof(float) : 0);
}

// Blocks (= partial sums) used for a batch of n samples.
int nhwc_bwd_weight_chunks(int n, int* spb_out) {
    int spb = ceil_div(n, 148);
    if (spb < 1) spb = 1;
    if (spb_out) *spb_out = spb;
    return ceil_div(n, spb);
}

// ... for a given layer.  A multi-channel layer with many output pixels (the pooled topology's 14x14 second convolution)
// has enough work per sample for a block of its own: twice the blocks, so two per SM hide each other's latencies.
int nhwc_bwd_weight_chunks_for(const b200cnn_conv_shape* s, int* spb_out) {
    if (s->c == 1 || s->oh * s->ow < 128) return nhwc_bwd_weight_chunks(s->n, spb_out);
    const int spb = std::max(1, ceil_div(s->n, 2 * 148));
    if (spb_out) *spb_out = spb;
    return ceil_div(s->n, spb);
}

// Floats of partial sums behind the 128-byte header that nhwc_conv_bwd_weight may use for this layer (whichever kernel
// it picks): what b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes and the engine allocate.
size_t nhwc_bwd_weight_scratch_floats(const b200cnn_conv_shape* s) {
    size_t need = size_t(nhwc_bwd_weight_chunks_for(s, nullptr)) * (size_t(s->oc) * s->c * s->k * s->k + s->oc);
    if (tc_conv3x3s2_bwd_weight_supported(s)) need = std::max(need, tc_conv3x3s2_bwd_weight_scratch_floats(s->n));
    if (tc_conv3x3s1_bwd_weight_supported(s)) need = std::max(need, tc_conv3x3s1_bwd_weight_scratch_floats(s->n));
    return need;
}

// Fused first layer of the pooled topology (see the kernels): conv 3x3 / stride 1 / pad 1 on one input channel, then a
// 2x2 / stride-2 max pool.
bool conv1_pool_supported(const b200cnn_conv_shape* s, int pool_k, int pool_stride) {
    return s->c == 1 && s->k == 3 && s->stride == 1 && s->pad == 1 && s->oh == s->h && s->ow == s->w && s->oh % 2 == 0 && s->ow % 2 == 0 &&
           pool_k == 2 && pool_stride == 2 && s->oc % 4 == 0 && s->oc >= 4 && kCpwThreads % s->oc == 0 && 10 * s->oc <= kCpwThreads &&
           size_t(s->h + 2) * (s->w + 2) * sizeof(float) <= 32 * 1024;
}

// x: [n][h][w] uint8 (scaled by 1/255) or float; yp / argmax: NHWC [n][oh/2][ow/2][oc]; a_full: NULL or NHWC [n][oh][ow][oc].
int conv1_pool_fwd(const void* x, int x_is_u8, const float* w, const float* b, float* a_full, float* yp, int32_t* argmax,
                   const b200cnn_conv_shape* s, int act, cudaStream_t st) {
    if (!conv1_pool_supported(s, 2, 2)) return set_error(kInvalidValue, "conv1_pool_fwd: unsupported shape");
    const size_t smem = ((size_t(s->h + 2) * (s->w + 2) + 3) / 4 * 4 + 10 * size_t(s->oc)) * sizeof(float);
    if (x_is_u8)
        B200CNN_CUDA(launch_kernel(conv1_pool_fwd_kernel<uint8_t>, dim3(s->n), dim3(kCpThreads), smem, st, static_cast<const uint8_t*>(x), w, b,
                                   a_full, yp, argmax, *s, act, 1.0f / 255.0f));
    else
        B200CNN_CUDA(launch_kernel(conv1_pool_fwd_kernel<float>, dim3(s->n), dim3(kCpThreads), smem, st, static_cast<const float*>(x), w, b,
                                   a_full, yp, argmax, *s, act, 1.0f));
    return 0;
}

// dyp / yp / argmax: the pool layer's gradient, output and winners (NHWC); act: the convolution's activation.
// partials: 128-byte header (completion counter) + chunks * (10 * oc) floats, chunks = nhwc_bwd_weight_chunks(n).
int conv1_pool_bwd_weight(const void* x, int x_is_u8, const float* dyp, const float* yp, const int32_t* argmax, int act, float* dw,
                          float* db, int accumulate, float* partials, unsigned* counter_in, const b200cnn_conv_shape* s, cudaStream_t st,
                          float* sgd_w, float* sgd_b, const float* lr_dev) {
    if (!conv1_pool_supported(s, 2, 2)) return set_error(kInvalidValue, "conv1_pool_bwd_weight: unsupported shape");
    unsigned* counter = counter_in ? counter_in : reinterpret_cast<unsigned*>(partials);
    partials += 32;
    int spb = 1;
    const int chunks = nhwc_bwd_weight_chunks(s->n, &spb);
    const size_t img = (size_t(s->h + 2) * (s->w + 2) + 3) / 4 * 4;
    const size_t smem = (size_t(spb) * img + size_t(kCpwThreads / s->oc) * 10 * s->oc) * sizeof(float);
    if (smem > 200 * 1024) return set_error(kInvalidValue, "conv1_pool_bwd_weight: samples do not fit shared memory");
    if (x_is_u8) {
        if (smem > 48 * 1024) cudaFuncSetAttribute(conv1_pool_bwd_weight_kernel<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        B200CNN_CUDA(launch_kernel(conv1_pool_bwd_weight_kernel<uint8_t>, dim3(chunks), dim3(kCpwThreads), smem, st, static_cast<const uint8_t*>(x),
                                   dyp, yp, argmax, act, partials, *s, spb, 1.0f / 255.0f, dw, db, accumulate, counter, sgd_w, sgd_b, lr_dev));
    } else {
        if (smem > 48 * 1024) cudaFuncSetAttribute(conv1_pool_bwd_weight_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        B200CNN_CUDA(launch_kernel(conv1_pool_bwd_weight_kernel<float>, dim3(chunks), dim3(kCpwThreads), smem, st, static_cast<const float*>(x),
                                   dyp, yp, argmax, act, partials, *s, spb, 1.0f, dw, db, accumulate, counter, sgd_w, sgd_b, lr_dev));
    }
    return 0;
}

// The tcgen05 weight-gradient kernel is used when the backend allows it (B200CNN_TC_WGRAD=0 keeps the CUDA-core
// kernel, for A/B timing).
static bool tc_wgrad_enabled() {
    static const bool env_on = [] {
        const char* v = getenv("B200CNN_TC_WGRAD");
        return !(v && v[0] == '0');
    }();
    return env_on && current_backend() != B200CNN_BACKEND_SIMT;
}

// x: NHWC float (or [n][h][w] uint8/float when c == 1); dz: NHWC, or CHW when dz_chw (c > 1 only).
// dw OIHW, db [oc].  partials: chunks*(nw+oc) floats.
// counter: one zero-initialised word used by the single-channel kernel's "last block reduces" step (it resets
// itself); NULL = the word right after this layer's partials.
int nhwc_conv_bwd_weight(const void* x, int x_is_u8, const float* dz, int dz_chw, float* dw, float* db, int accumulate,
                         float* partials, unsigned* counter_in, const b200cnn_conv_shape* s, cudaStream_t st, float* sgd_w,
                         float* sgd_b, const float* lr_dev, const TailXchg* xchg, long w_off, long b_off) {
    if (xchg && !(s->c == 1 && x_is_u8 && tc_conv1_bwd_weight_enabled()))
        return set_error(kInvalidValue, "nhwc_conv_bwd_weight: the in-kernel all-reduce exists only in the tensor-core conv1 kernel");
    if (sgd_w && s->c != 1) return set_error(kInvalidValue, "nhwc_conv_bwd_weight: the fused update exists only in the single-channel kernel");
    if (!nhwc_bwd_weight_supported(s, x_is_u8)) return set_error(kInvalidValue, "nhwc_conv_bwd_weight: unsupported shape");
    // Scratch layout: [128-byte header | partial sums].  The header's first word is the completion counter of the
    // single-channel kernel; it resets itself, and no kernel writes partial sums over it, so one scratch buffer can
    // be shared by the layers of a network as long as it was zeroed once.
    unsigned* header = reinterpret_cast<unsigned*>(partials);
    partials += 32;
    int spb = 1;
    const int chunks = nhwc_bwd_weight_chunks_for(s, &spb);
    const int kk = s->k * s->k, nw = s->oc * s->c * kk;
    if (s->c == 1) {
        if (dz_chw) return set_error(kInvalidValue, "nhwc_conv_bwd_weight: CHW gradients are only read by the multi-channel kernel");
        const int nout = nw + s->oc;
        const int items = (s->oc / 4) * (kk + 1);
        if (items > kC1Threads || nout > kC1Threads) return set_error(kInvalidValue, "nhwc_conv_bwd_weight: layer too large for the single-channel kernel");
        const int slices = kC1Threads / items;
        unsigned* counter = counter_in ? counter_in : header;
        if (x_is_u8 && tc_conv1_bwd_weight_enabled() && tc_conv1_bwd_weight_supported(s, chunks))
            return tc_conv1_bwd_weight(static_cast<const uint8_t*>(x), dz, dw, db, accumulate, partials, counter, s, st, sgd_w, sgd_b,
                                       lr_dev, xchg, w_off, b_off);
        if (xchg) return set_error(kInvalidValue, "nhwc_conv_bwd_weight: shape not supported by the tensor-core conv1 kernel");
        const size_t smem = (size_t(spb) * (((s->h * s->w + 3) & ~3) + size_t(s->oh) * s->ow * s->oc) + size_t(std::max(slices, kC1Threads / nout)) * nout) * sizeof(float);
        if (smem > 200 * 1024) return set_error(kInvalidValue, "nhwc_conv_bwd_weight: samples do not fit shared memory");
        if (x_is_u8) {
            if (smem > 48 * 1024) cudaFuncSetAttribute(conv_bwd_weight_c1_nhwc_kernel<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            B200CNN_CUDA(launch_kernel(conv_bwd_weight_c1_nhwc_kernel<uint8_t>, dim3(chunks), dim3(kC1Threads), smem, st, static_cast<const uint8_t*>(x), dz, partials, *s, spb, 1.0f / 255.0f, dw, db, accumulate, counter, sgd_w, sgd_b, lr_dev));
        } else {
            if (smem > 48 * 1024) cudaFuncSetAttribute(conv_bwd_weight_c1_nhwc_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            B200CNN_CUDA(launch_kernel(conv_bwd_weight_c1_nhwc_kernel<float>, dim3(chunks), dim3(kC1Threads), smem, st, static_cast<const float*>(x), dz, partials, *s, spb, 1.0f, dw, db, accumulate, counter, sgd_w, sgd_b, lr_dev));
        }
    } else if (tc_wgrad_enabled() && !x_is_u8 && tc_conv3x3s2_bwd_weight_supported(s)) {
        // tensor cores: split-K implicit GEMM + a small reduce kernel over the n/8 partial tiles
        return tc_conv3x3s2_bwd_weight(static_cast<const float*>(x), dz, dz_chw, dw, db, accumulate, partials, s, st);
    } else if (tc_wgrad_enabled() && !x_is_u8 && !dz_chw && tc_conv3x3s1_bwd_weight_supported(s)) {
        // the same split-K implicit GEMM over the stride-1 views (pooled topology)
        return tc_conv3x3s1_bwd_weight(static_cast<const float*>(x), dz, dw, db, accumulate, partials, s, st);
    } else {
        const int work = (s->oc / 4) * (s->c / 4) * kk;
        const int threads = ((work + 31) / 32) * 32;
        const size_t smem = size_t(spb) * (size_t(s->h) * s->w * s->c + size_t(s->oh) * s->ow * s->oc) * sizeof(float);
        if (smem > 200 * 1024) return set_error(kInvalidValue, "nhwc_conv_bwd_weight: samples do not fit shared memory");
        if (smem > 48 * 1024) cudaFuncSetAttribute(conv_bwd_weight_nhwc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        B200CNN_CUDA(launch_kernel(conv_bwd_weight_nhwc_kernel, dim3(chunks), dim3(threads), smem, st, static_cast<const float*>(x), dz, partials, *s, spb, dz_chw));
    }
    B200CNN_CHECK_LAUNCH("nhwc_conv_bwd_weight");
    if (s->c == 1) return 0;  // reduced by its last block
    return reduce_conv_partials(partials, dw, db, chunks, nw, s->oc, accumulate, st);
}

}  // namespace b200cnn
