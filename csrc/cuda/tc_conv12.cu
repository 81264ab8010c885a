// Forward of BOTH convolutions of the reference topology as one kernel — the ROOT of the captured step.
//     a1 = relu(conv3x3/s2(x / 255, W1) + b1)      1 -> 16 channels, 28x28 -> 14x14   (CUDA cores, 9 FMAs per output)
//     a2 = relu(conv3x3/s2(a1, W2) + b2)           16 -> 32 channels, 14x14 -> 7x7    (tcgen05 implicit GEMM)
// Why fuse exactly these two: the second kernel node of a launched CUDA graph does not start before ~6.2 us after the
// root started, whatever the root does (measured, engine.cu: B200CNN_GRAPH_ROOT).  With conv1 as the root (4.5 us) the
// tensor-core convolution began at 6.1 us and delivered a2 at 13 us.  Here a CTA owns two images: all of its threads
// compute a1 from the uint8 images (shared memory) and write every pixel both to global memory (NHWC, backward needs
// it) and straight into the K-major, 64-byte-swizzled A operand tiles of the second convolution — pixel (y, x) is row
// (j, i) of tap (ky, kx) wherever y = 2j - 1 + ky, x = 2i - 1 + kx: 1, 2 or 4 tiles — so no TMA load, no kernel
// boundary and no global round trip separates the two layers, and the root kernel fills the launch ramp with the work
// of the second node.
//
// Replaces Layer_feedForw_conv for layers 1 and 2 (/root/reference/cnn.c:175-211) and the reference's only kernel,
// conv_forward_kernel (/root/reference/CUDAcnn.cu:167-195).
#include <cstdlib>

#include "backend.cuh"
#include "common.cuh"
#include "tc_common.cuh"

namespace b200cnn {

namespace {

using namespace tc;

constexpr int kThreads = 384;  // all: operand loads + conv1; warp 1: TMEM + MMA; warps 2-5: epilogue
constexpr int kTaps = 9, kTileRows = 128;
constexpr int kC1 = 16, kOC2 = 32;       // channels after conv1 / conv2
constexpr int kImgs = 2, kIn = 28, kMid = 14, kOut = 7, kPix2 = kOut * kOut;
constexpr int kATile = kTileRows * kC1 * 4;  // 8 KiB per tap, 64-byte rows
constexpr int kBTile = kOC2 * kC1 * 4;       // 2 KiB per tap
constexpr int kOffB = kTaps * kATile;
constexpr int kOffXs = kOffB + kTaps * kBTile;          // uint8 images
constexpr int kOffW1 = kOffXs + kImgs * kIn * kIn;      // [9][16] floats + [16] bias
constexpr int kOffBars = kOffW1 + (kTaps * kC1 + kC1) * 4;
constexpr size_t kSmem = size_t(kOffBars) + 32 + 1024;
static_assert(kOffXs % 16 == 0 && kOffW1 % 16 == 0 && kOffBars % 8 == 0, "alignment");

struct Conv12Args {
    const uint8_t* x;   // [n][28][28]
    const float* w1;    // OIHW [16][1][3][3]
    const float* b1;    // [16]
    float* a1;          // NHWC [n][14][14][16]
    const float* w2;    // OIHW [32][16][3][3]
    const float* b2;    // [32]
    float* y;           // [n][ldy], CHW-flattened per sample
    int ldy, n, act1, act2;
};

__global__ void __launch_bounds__(kThreads, 1)
tc_conv12_fwd_kernel(Conv12Args a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_tiles = smem;
    uint8_t* b_tiles = smem + kOffB;
    uint8_t* xs = smem + kOffXs;
    float* w1s = reinterpret_cast<float*>(smem + kOffW1);  // [tap][oc]
    float* b1s = w1s + kTaps * kC1;
    uint64_t* tmem_full_bar = reinterpret_cast<uint64_t*>(smem + kOffBars);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int t = threadIdx.x, warp = t / 32, lane = t % 32;
    const int n0 = blockIdx.x * kImgs;
    const int ns = min(kImgs, a.n - n0);
    const int rows = ns * kPix2;

    if (warp == 0 && lane == 0) {
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<32>(tmem_slot);
    // Tile rows whose source pixel lies in the zero padding (top row of the ky = 0 taps, left column of the kx = 0 taps)
    for (int i = t; i < kTaps * kImgs * kPix2; i += kThreads) {
        const int tap = i / (kImgs * kPix2), row = i % (kImgs * kPix2);
        const int p = row % kPix2, j = p / kOut, ii = p % kOut;
        if (2 * j - 1 + tap / 3 < 0 || 2 * ii - 1 + tap % 3 < 0) {
            float4* d = reinterpret_cast<float4*>(a_tiles + tap * kATile + row * 64);
#pragma unroll
            for (int c = 0; c < 4; c++) d[c] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();

    // ---- operands from global memory: W2, images, W1 / b1 — EVERY load of a thread is issued before the first use, so the
    //      whole phase is one L2 round trip ----
    {
        constexpr int kW = kOC2 * kC1 * kTaps, kBatch = kW / kThreads;  // 12 weights of W2 per thread
        static_assert(kW % kThreads == 0, "W2 splits evenly over the threads");
        float v[kBatch];
#pragma unroll
        for (int u = 0; u < kBatch; u++) v[u] = __ldg(a.w2 + t + u * kThreads);
        uchar4 px = make_uchar4(0, 0, 0, 0);
        const bool has_px = t < ns * kIn * kIn / 4;  // 392 uchar4 for two images
        if (has_px) px = __ldg(reinterpret_cast<const uchar4*>(a.x + size_t(n0) * kIn * kIn) + t);
        const uchar4 px2 = (t + kThreads < ns * kIn * kIn / 4) ? __ldg(reinterpret_cast<const uchar4*>(a.x + size_t(n0) * kIn * kIn) + t + kThreads)
                                                                : make_uchar4(0, 0, 0, 0);
        const float w1v = t < kTaps * kC1 ? __ldg(a.w1 + t) : 0.f, b1v = t < kC1 ? __ldg(a.b1 + t) : 0.f;
        if (has_px) reinterpret_cast<uchar4*>(xs)[t] = px;
        if (t + kThreads < ns * kIn * kIn / 4) reinterpret_cast<uchar4*>(xs)[t + kThreads] = px2;
        if (t < kTaps * kC1) w1s[(t % kTaps) * kC1 + t / kTaps] = w1v;  // OIHW -> [tap][oc]
        if (t < kC1) b1s[t] = b1v;
        // W2: OIHW -> per-tap K-major [oc][c] tiles, 64-byte swizzle (16-byte chunk ^= (row >> 1) & 3)
#pragma unroll
        for (int u = 0; u < kBatch; u++) {
            const int i = t + u * kThreads;
            const int tap = i % kTaps, c = (i / kTaps) % kC1, oc = i / (kTaps * kC1);
            const int chunk = (c >> 2) ^ ((oc >> 1) & 3);
            *reinterpret_cast<float*>(b_tiles + tap * kBTile + oc * 64 + chunk * 16 + (c & 3) * 4) = v[u];
        }
    }
    __syncthreads();

    // ---- conv1 on the CUDA cores: item = (pixel of a1, half of its 16 channels) ----
    for (int item = t; item < ns * kMid * kMid * 2; item += kThreads) {
        const int h = item & 1, pp = item >> 1;
        const int g = pp / (kMid * kMid), p = pp % (kMid * kMid), y = p / kMid, x = p % kMid;
        float acc[8];
#pragma unroll
        for (int c = 0; c < 8; c++) acc[c] = b1s[8 * h + c];
        const uint8_t* img = xs + g * kIn * kIn;
#pragma unroll
        for (int ky = 0; ky < 3; ky++) {
            const int iy = 2 * y - 1 + ky;
            if (iy < 0) continue;  // (iy <= 27 always)
#pragma unroll
            for (int kx = 0; kx < 3; kx++) {
                const int ix = 2 * x - 1 + kx;
                if (ix < 0) continue;
                const float xv = float(img[iy * kIn + ix]) * (1.0f / 255.0f);
                const float4 wa = *reinterpret_cast<const float4*>(w1s + (ky * 3 + kx) * kC1 + 8 * h);
                const float4 wb = *reinterpret_cast<const float4*>(w1s + (ky * 3 + kx) * kC1 + 8 * h + 4);
                acc[0] = fmaf(xv, wa.x, acc[0]); acc[1] = fmaf(xv, wa.y, acc[1]);
                acc[2] = fmaf(xv, wa.z, acc[2]); acc[3] = fmaf(xv, wa.w, acc[3]);
                acc[4] = fmaf(xv, wb.x, acc[4]); acc[5] = fmaf(xv, wb.y, acc[5]);
                acc[6] = fmaf(xv, wb.z, acc[6]); acc[7] = fmaf(xv, wb.w, acc[7]);
            }
        }
        const float4 lo = make_float4(apply_act(a.act1, acc[0]), apply_act(a.act1, acc[1]), apply_act(a.act1, acc[2]), apply_act(a.act1, acc[3]));
        const float4 hi = make_float4(apply_act(a.act1, acc[4]), apply_act(a.act1, acc[5]), apply_act(a.act1, acc[6]), apply_act(a.act1, acc[7]));
        float4* gp = reinterpret_cast<float4*>(a.a1 + (size_t(n0 + g) * kMid * kMid + p) * kC1 + 8 * h);
        gp[0] = lo;
        gp[1] = hi;
        // ... and into every tap tile that reads this pixel: y = 2j - 1 + ky, x = 2i - 1 + kx
#pragma unroll
        for (int ky = 0; ky < 3; ky++) {
            const int jy = y + 1 - ky;
            if ((jy & 1) || jy < 0 || jy >= 2 * kOut) continue;
#pragma unroll
            for (int kx = 0; kx < 3; kx++) {
                const int ix = x + 1 - kx;
                if ((ix & 1) || ix < 0 || ix >= 2 * kOut) continue;
                const int row = g * kPix2 + (jy >> 1) * kOut + (ix >> 1);
                uint8_t* d = a_tiles + (ky * 3 + kx) * kATile + row * 64;
                const int sw = (row >> 1) & 3;
                *reinterpret_cast<float4*>(d + (((2 * h) ^ sw) << 4)) = lo;
                *reinterpret_cast<float4*>(d + (((2 * h + 1) ^ sw) << 4)) = hi;
            }
        }
    }
    fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor-core (async) proxy
    __syncthreads();

    if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(kTileRows, kOC2, false, false);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_tiles), b_addr = smem_u32(b_tiles);
#pragma unroll
            for (int tap = 0; tap < kTaps; tap++)
#pragma unroll
                for (int kk = 0; kk < kC1 / 8; kk++)
                    umma_tf32(tmem_base, make_smem_desc(a_addr + tap * kATile + kk * 32, 16, 512, kSw64),
                              make_smem_desc(b_addr + tap * kBTile + kk * 32, 16, 512, kSw64), idesc, (tap | kk) != 0);
            umma_commit(tmem_full_bar);
        }
    } else if (warp >= 2 && warp < 6) {
        const int q = warp & 3;
        const int r = 32 * q + lane;
        float bias[kOC2];  // fetched before the accumulator is ready
#pragma unroll
        for (int i = 0; i < kOC2; i++) bias[i] = __ldg(a.b2 + i);
        mbar_wait(tmem_full_bar, 0);
        tcgen05_fence_after();
        const bool live = r < rows;
        const int g = r / kPix2, p = r % kPix2;
        float* out = a.y + size_t(n0 + g) * a.ldy + p;
#pragma unroll
        for (int c0 = 0; c0 < kOC2; c0 += 16) {
            float v[16];
            __syncwarp();
            tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(c0), v);
            if (live) {
#pragma unroll
                for (int i = 0; i < 16; i++) out[size_t(c0 + i) * kPix2] = apply_act(a.act2, v[i] + bias[c0 + i]);
            }
        }
        __syncwarp();
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<32>(tmem_base);
    }
}

}  // namespace

bool tc_conv12_fwd_enabled() {
    static const bool on = [] {
        const char* v = getenv("B200CNN_FUSED_C12");  // B200CNN_FUSED_C12=0: the two forward convolutions as two kernels (A/B timing)
        return !(v && v[0] == '0');
    }();
    return on && current_backend() != B200CNN_BACKEND_SIMT;
}

bool tc_conv12_fwd_supported(const b200cnn_conv_shape* s1, const b200cnn_conv_shape* s2) {
    return s1->c == 1 && s1->oc == kC1 && s1->k == 3 && s1->stride == 2 && s1->pad == 1 && s1->h == kIn && s1->w == kIn && s1->oh == kMid &&
           s1->ow == kMid && s2->c == kC1 && s2->oc == kOC2 && s2->k == 3 && s2->stride == 2 && s2->pad == 1 && s2->h == kMid && s2->w == kMid &&
           s2->oh == kOut && s2->ow == kOut && s1->n == s2->n && s1->n >= 1;
}

// x: uint8 [n][28][28]; a1: NHWC [n][14][14][16]; y: [n][ldy] with the sample's 32 x 7 x 7 outputs CHW-flattened.
int tc_conv12_fwd(const uint8_t* x, const float* w1, const float* b1, float* a1_nhwc, int act1, const float* w2, const float* b2, float* y,
                  int ldy, int act2, const b200cnn_conv_shape* s1, const b200cnn_conv_shape* s2, cudaStream_t st) {
    if (!tc_conv12_fwd_supported(s1, s2)) return set_error(kInvalidValue, "tc_conv12_fwd: unsupported shapes");
    Conv12Args a{x, w1, b1, a1_nhwc, w2, b2, y, ldy, s1->n, act1, act2};
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(tc_conv12_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kSmem)));
        configured = true;
    }
    B200CNN_CUDA(launch_kernel(tc_conv12_fwd_kernel, dim3(ceil_div(s1->n, kImgs)), dim3(kThreads), kSmem, st, a));
    B200CNN_CHECK_LAUNCH("tc_conv12_fwd");
    return 0;
}

}  // namespace b200cnn
