// The tail of the backward pass of the reference topology as ONE kernel:
//     dz1 = relu'(a1) * conv_transpose(dz2, W2)           (second convolution's data gradient, tc_conv.cu)
//     dW1 = sum_pixels dz1 (x) patches(x),  db1 = sum dz1   (first convolution's weight gradient, tc_conv1_wgrad.cu)
//     + the last-block reduction over the CTAs, the fused SGD step / the in-kernel all-reduce with the peers.
// Both kernels already gave a CTA the same two images, and the thread that owns an output pixel (j, i) of the second
// convolution after tcgen05.ld holds dz1 of the four input pixels (2j+py, 2i+px) x 16 channels — which is exactly two
// rows of the first convolution's M-major A operand ([pixel pair (y, i)][(parity px, oc)], 128 bytes).  So instead of
// storing dz1 to global memory, ending the kernel, launching the next one and loading dz1 back through TMA, the
// epilogue threads write their rows straight into the (swizzled) operand tile in shared memory — it reuses the space of
// the gradient tiles the first GEMM is done with — and the same MMA thread issues the second GEMM (25 x M=64, N=32,
// K=8) against the tap tile that warps 6-11 built from the uint8 images before griddepcontrol.wait.  dz1 never exists
// in global memory; one kernel boundary, one TMA round trip and 3.2 MB of stores + loads leave the critical path.
// MEASURED SLOWER in the step (see tc_conv_tail_enabled below), hence opt-in.
//
// Replaces the data half of Layer_feedBack_conv for layer 2 and the weight half for layer 1
// (/root/reference/cnn.c:213-247) plus the update of layer 1 (cnn.c:249-262 / cnnmpi.c:490).
#include <cstdlib>

#include "../comm/xchg_device.cuh"
#include "backend.cuh"
#include "common.cuh"
#include "tc_common.cuh"

namespace b200cnn {

namespace {

using namespace tc;

constexpr int kThreads = 384;   // warp 0: TMA, warp 1: TMEM + both MMA loops, warps 2-5: weights + epilogues, warps 6-11: tap tile; all: tail
constexpr int kP1 = 192;        // threads of the first GEMM's CHW repack (the mapping of tc_conv3x3s2_bwd_data_kernel)
constexpr int kTaps = 9, kTileRows = 128, OC = 32, C = 16;
constexpr int kATile = kTileRows * OC * 4;  // 16 KiB per shift (dj, di)
constexpr int kBTile = C * OC * 4;          // [c rows][oc] per tap
constexpr int kImgs = 2, kPix2 = 49, kPairs = 98, kRows1 = kImgs * kPairs, kSteps1 = (kRows1 + 7) / 8, kTile1 = kSteps1 * 8 * 128;
constexpr int kOC1 = 16, kNW = kOC1 * kTaps, kNOut = kNW + kOC1, kInPix = 28 * 28, kParts = kThreads / kNOut;
constexpr int kBuilders = kThreads - kP1;   // 192
// shared memory map (bytes from the 1 KiB-aligned base)
constexpr int kOffB = 4 * kATile;                       // 9 tap tiles of W2
constexpr int kOffStage = kOffB + kTaps * kBTile;       // CHW staging of dz2
constexpr int kOffB1 = kOffStage + kTileRows * OC * 4;  // tap tile of the first convolution (multiple of 1 KiB)
constexpr int kOffXs = kOffB1 + kTile1;                 // the CTA's two images
constexpr int kOffPart = kOffXs + kImgs * kInPix;
constexpr int kOffRed = kOffPart + 2 * kOC1 * 10 * 4;
constexpr int kOffBars = kOffRed + kParts * kNOut * 4;
constexpr size_t kSmem = size_t(kOffBars) + 64 + 1024;
static_assert(kOffB1 % 1024 == 0 && kOffXs % 16 == 0 && kOffBars % 8 == 0, "alignment");
static_assert(kTile1 + kTile1 <= 4 * kATile, "the A operand of the second GEMM (and the group behind it that M=64 reads) lives in the dead dz2 tiles");

struct TailArgs {
    const float* dz_chw;   // dz2 as [n][32][7][7] (what the FC layer writes), or NULL: NHWC behind tmap_dz
    const float* w2;       // OIHW [32][16][3][3]
    const float* a_prev;   // NHWC [n][14][14][16] activations of layer 1 (ReLU mask) or NULL
    float* dx;             // NULL (nobody reads dz1) or NHWC [n][14][14][16]
    const uint8_t* x;      // [n][28][28]
    float* partials;       // [gridDim.x][160]
    float* dw;             // [16][9]
    float* db;             // [16]
    unsigned* counter;     // zero before the first launch; wraps back to zero
    float* sgd_w;          // optional fused update of layer 1
    float* sgd_b;
    const float* lr_dev;
    int n, act_prev, early_w;
    TailXchg xc;           // data parallel: the last block all-reduces the 160 numbers with the peers (tc_conv1_wgrad.cu)
    long w_off, b_off;
};

__global__ void __launch_bounds__(kThreads, 1)
tc_conv2dx_conv1dw_kernel(const __grid_constant__ CUtensorMap tmap_dz, TailArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_tiles = smem;              // 4 shifted views of dz2; later: [pixel pairs][(px, oc)] of dz1
    uint8_t* b_tiles = smem + kOffB;
    float* stage = reinterpret_cast<float*>(smem + kOffStage);
    uint8_t* a1_tile = smem;
    uint8_t* b1_tile = smem + kOffB1;
    uint8_t* xs = smem + kOffXs;
    float* part = reinterpret_cast<float*>(smem + kOffPart);  // [px][oc][10]
    float* red = reinterpret_cast<float*>(smem + kOffRed);    // [kParts][160]
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kOffBars);
    uint64_t* tmem_full_bar = full_bar + 1;  // first GEMM complete
    uint64_t* a1_bar = full_bar + 2;         // dz1 rows are in the operand tile (128 epilogue threads)
    uint64_t* tmem2_bar = full_bar + 3;      // second GEMM complete
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(full_bar + 4);
    __shared__ unsigned is_last;

    const int t = threadIdx.x, warp = t / 32, lane = t % 32;
    const int n0 = blockIdx.x * kImgs;
    const int ns = min(kImgs, a.n - n0);
    const int rows = ns * kPix2;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_dz);
        mbar_init(full_bar, 1);
        mbar_init(tmem_full_bar, 1);
        mbar_init(a1_bar, 128);
        mbar_init(tmem2_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<128>(tmem_slot);
    // W2 per tap as the B operand [n = c][k = oc], K-major 128-byte rows, 128B swizzle (16-byte chunk ^= row & 7)
    auto gather_weights = [&]() {
        constexpr int kW = OC * C * kTaps, kStride = 128, kBatch = 12;
        for (int i0 = t - 64; i0 < kW; i0 += kBatch * kStride) {
            float v[kBatch];
#pragma unroll
            for (int u = 0; u < kBatch; u++) v[u] = (i0 + u * kStride < kW) ? __ldg(a.w2 + i0 + u * kStride) : 0.f;
#pragma unroll
            for (int u = 0; u < kBatch; u++) {
                const int i = i0 + u * kStride;
                if (i >= kW) break;
                const int tp = i % kTaps, c = (i / kTaps) % C, oc = i / (kTaps * C);
                const int chunk = (oc >> 2) ^ (c & 7);
                *reinterpret_cast<float*>(b_tiles + tp * kBTile + c * 128 + chunk * 16 + (oc & 3) * 4) = v[u];
            }
        }
    };
    if (a.early_w && warp >= 2 && warp < 6) gather_weights();
    if (warp >= 6) {
        // ---- tap tile of the first convolution: [pixel pair q][(px', j)], independent of every producer ----
        const int tb = t - kP1;
        for (int i = tb; i < ns * kInPix / 4; i += kBuilders)
            reinterpret_cast<uchar4*>(xs)[i] = __ldg(reinterpret_cast<const uchar4*>(a.x + size_t(n0) * kInPix) + i);
        asm volatile("bar.sync 1, %0;" ::"n"(kBuilders) : "memory");
        const int e = lane >> 4, j = lane & 15;
        const int ky = j / 3, kx = j % 3;
        for (int r = warp - 6; r < kRows1 + 4; r += kBuilders / 32) {
            float v = 0.0f;
            const int g = r / kPairs, q = r % kPairs;
            if (r < kRows1 && g < ns) {
                if (j < kTaps) {
                    const int y = q / 7, px = 2 * (q % 7) + e;
                    const int iy = 2 * y - 1 + ky, ix = 2 * px - 1 + kx;
                    if (iy >= 0 && iy < 28 && ix >= 0 && ix < 28) v = float(xs[g * kInPix + iy * 28 + ix]) * (1.0f / 255.0f);
                } else if (j == kTaps) {
                    v = 1.0f;
                }
            }
            *reinterpret_cast<float*>(b1_tile + r * 128 + ((((lane >> 3) ^ (r & 3))) << 5) + (lane & 7) * 4) = v;
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();

    if (a.dz_chw) {
        // dz2 arrives CHW: coalesced float4 loads into the staging buffer, then a shared -> shared transpose into the four
        // shifted, swizzled K-major tiles (same mapping as tc_conv3x3s2_bwd_data_kernel, threads 0..191).
        const int nfl = rows * OC;
        {
            const float4* src = reinterpret_cast<const float4*>(a.dz_chw + size_t(n0) * OC * kPix2);
            for (int i = t; i < nfl / 4; i += kThreads) reinterpret_cast<float4*>(stage)[i] = __ldg(src + i);
        }
        __syncthreads();
        if (t < kP1) {
            const int r = t % kTileRows, half = t / kTileRows;
            if (r < rows) {
                const int g = r / kPix2, p = r % kPix2, j = p / 7, i = p % 7;
                const float* base = stage + g * OC * kPix2;
                for (int sft = 0; sft < 4; sft++) {
                    const int jj = j + (sft >> 1), ix = i + (sft & 1);
                    const bool inb = jj < 7 && ix < 7;
                    const float* src = base + jj * 7 + ix;
                    uint8_t* dst = a_tiles + sft * kATile + r * 128;
                    const int lo = (r < kP1 - kTileRows) ? half * (OC / 8) : 0;
                    const int hi = (r < kP1 - kTileRows) ? lo + OC / 8 : (half == 0 ? OC / 4 : 0);
                    for (int oc4 = lo; oc4 < hi; oc4++) {
                        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (inb) {
                            const float* s4 = src + oc4 * 4 * kPix2;
                            v = make_float4(s4[0], s4[kPix2], s4[2 * kPix2], s4[3 * kPix2]);
                        }
                        *reinterpret_cast<float4*>(dst + ((oc4 ^ (r & 7)) << 4)) = v;
                    }
                }
            }
        }
    } else if (warp == 0 && lane == 0) {
        mbar_arrive_expect_tx(full_bar, uint32_t(4 * kImgs * kPix2 * OC * 4));
#pragma unroll
        for (int sft = 0; sft < 4; sft++) tma_load_4d(a_tiles + sft * kATile, &tmap_dz, full_bar, 0, sft & 1, sft >> 1, n0);
    }
    if (!a.early_w && warp >= 2 && warp < 6) gather_weights();
    fence_proxy_async_smem();
    __syncthreads();

    if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(kTileRows, C, false, false);
            if (!a.dz_chw) mbar_wait(full_bar, 0);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_tiles), b_addr = smem_u32(b_tiles);
            // input pixel (2j+py, 2i+px) receives tap (ky, kx) from output pixel (j+dj, i+di)
#pragma unroll
            for (int cls = 0; cls < 4; cls++) {
                const int py = cls >> 1, px = cls & 1;
                bool first = true;
#pragma unroll
                for (int ky = 0; ky < 3; ky++) {
                    if ((py == 0) != (ky == 1)) continue;
                    const int dj = (ky == 0) ? 1 : 0;
#pragma unroll
                    for (int kx = 0; kx < 3; kx++) {
                        if ((px == 0) != (kx == 1)) continue;
                        const int di = (kx == 0) ? 1 : 0;
                        const int sft = dj * 2 + di, tp = ky * 3 + kx;
#pragma unroll
                        for (int kk = 0; kk < OC / 8; kk++)
                            umma_tf32(tmem_base + cls * C, make_smem_desc(a_addr + sft * kATile + kk * 32, 16, 1024, kSw128),
                                      make_smem_desc(b_addr + tp * kBTile + kk * 32, 16, 1024, kSw128), idesc, !(first && kk == 0));
                        first = false;
                    }
                }
            }
            umma_commit(tmem_full_bar);
            // ---- second GEMM: D[(px, oc)][(px', j)] = sum over pixel pairs of dz1 x taps ----
            constexpr uint32_t idesc1 = make_idesc_tf32(64, 32, true, true);
            mbar_wait(a1_bar, 0);
            tcgen05_fence_after();
            const uint32_t a1_addr = smem_u32(a1_tile), b1_addr = smem_u32(b1_tile);
#pragma unroll 5
            for (int kk = 0; kk < kSteps1; kk++)
                umma_tf32(tmem_base + 64, make_smem_desc(a1_addr + kk * 1024, kTile1, 512, kSw128Base32),
                          make_smem_desc(b1_addr + kk * 1024, 4096, 512, kSw128Base32), idesc1, kk != 0);
            umma_commit(tmem2_bar);
        }
    } else if (warp >= 2 && warp < 6) {
        const int q = warp & 3;
        const int r = 32 * q + lane;
        const bool live = r < rows;
        const int g = r / kPix2, p = r % kPix2, j = p / 7, i = p % 7;
        float4 msk[4][C / 4];
        if (live && a.a_prev) {
#pragma unroll
            for (int cls = 0; cls < 4; cls++) {
                const int y = 2 * j + (cls >> 1), x = 2 * i + (cls & 1);
                const float4* ap = reinterpret_cast<const float4*>(a.a_prev + ((size_t(n0 + g) * 14 + y) * 14 + x) * C);
#pragma unroll
                for (int k4 = 0; k4 < C / 4; k4++) msk[cls][k4] = __ldg(ap + k4);
            }
        }
        mbar_wait(tmem_full_bar, 0);  // every MMA of the first GEMM is done: the dz2 tiles are dead
        tcgen05_fence_after();
#pragma unroll
        for (int cls = 0; cls < 4; cls++) {
            float v[16];
            __syncwarp();
            tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(cls * C), v);
            if (r >= kImgs * kPix2) continue;  // rows 98..127 of the tile are nobody's pixel
            const int py = cls >> 1, px = cls & 1;
            if (live) {
                if (a.a_prev) {
#pragma unroll
                    for (int k4 = 0; k4 < C / 4; k4++) {
                        const float4 av = msk[cls][k4];
                        v[4 * k4 + 0] *= act_grad_from_output(a.act_prev, av.x);
                        v[4 * k4 + 1] *= act_grad_from_output(a.act_prev, av.y);
                        v[4 * k4 + 2] *= act_grad_from_output(a.act_prev, av.z);
                        v[4 * k4 + 3] *= act_grad_from_output(a.act_prev, av.w);
                    }
                }
                if (a.dx) {
                    float4* dp = reinterpret_cast<float4*>(a.dx + ((size_t(n0 + g) * 14 + 2 * j + py) * 14 + 2 * i + px) * C);
#pragma unroll
                    for (int k4 = 0; k4 < C / 4; k4++) dp[k4] = make_float4(v[4 * k4], v[4 * k4 + 1], v[4 * k4 + 2], v[4 * k4 + 3]);
                }
            } else {
#pragma unroll
                for (int k = 0; k < 16; k++) v[k] = 0.0f;  // an image past the end of the batch
            }
            // row (g, y = 2j+py, pair i) of the M-major operand: this pixel is its half px; 32-byte chunk ^= row & 3
            const int ra = g * kPairs + (2 * j + py) * 7 + i;
            uint8_t* row = a1_tile + ra * 128;
#pragma unroll
            for (int h = 0; h < 2; h++) {
                float4* d4 = reinterpret_cast<float4*>(row + (((2 * px + h) ^ (ra & 3)) << 5));
                d4[0] = make_float4(v[8 * h], v[8 * h + 1], v[8 * h + 2], v[8 * h + 3]);
                d4[1] = make_float4(v[8 * h + 4], v[8 * h + 5], v[8 * h + 6], v[8 * h + 7]);
            }
        }
        if (r >= 124) {  // the four K rows behind the last pixel pair
            float4* z = reinterpret_cast<float4*>(a1_tile + (kRows1 + r - 124) * 128);
#pragma unroll
            for (int k = 0; k < 8; k++) z[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        fence_proxy_async_smem();
        mbar_arrive(a1_bar);
        if (warp == 4 || warp == 5) {
            // TMEM lane quarters 0 and 1 hold rows (px = 0, oc) and (px = 1, oc) in their first 16 lanes (M=64 layout)
            const int e = warp & 3;
            mbar_wait(tmem2_bar, 0);
            tcgen05_fence_after();
            float lo[16], hi[16];
            tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * e) << 16) + 64u, lo);
            tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * e) << 16) + 80u, hi);
            if (lane < 16) {
#pragma unroll
                for (int jj = 0; jj < 10; jj++) part[(e * kOC1 + lane) * 10 + jj] = e == 0 ? lo[jj] : hi[jj];
            }
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<128>(tmem_base);
    }
    if (t < kNOut) {
        const int oc = t < kNW ? t / kTaps : t - kNW, jj = t < kNW ? t % kTaps : kTaps;
        a.partials[size_t(blockIdx.x) * kNOut + t] = part[oc * 10 + jj] + part[(kOC1 + oc) * 10 + jj];
    }
    // ---- the block that finishes last adds all partials in block order (same tail as tc_conv1_bwd_weight_kernel) ----
    __syncthreads();
    if (t == 0) {
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
        const unsigned done = atomicInc(a.counter, gridDim.x - 1);
        is_last = (done == gridDim.x - 1);
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
    __syncthreads();
    if (!is_last) return;
    const int jo = t % kNOut, pt = t / kNOut;
    if (pt < kParts) {
        float acc = 0.f;
        int c = pt;
        for (; c + 31 * kParts < int(gridDim.x); c += 32 * kParts) {
            float v[32];
#pragma unroll
            for (int u = 0; u < 32; u++) v[u] = __ldcg(a.partials + size_t(c + u * kParts) * kNOut + jo);
#pragma unroll
            for (int u = 0; u < 32; u++) acc += v[u];
        }
        for (; c < int(gridDim.x); c += kParts) acc += __ldcg(a.partials + size_t(c) * kNOut + jo);
        red[pt * kNOut + jo] = acc;
    }
    __syncthreads();
    float acc = 0.f;
    if (t < kNOut) {
        acc = red[t];
        for (int sl = 1; sl < kParts; sl++) acc += red[sl * kNOut + t];
        float* dst = t < kNW ? (a.dw ? a.dw + t : nullptr) : (a.db ? a.db + (t - kNW) : nullptr);
        if (dst) *dst = acc;
        if (a.sgd_w) {
            float* wdst = t < kNW ? a.sgd_w + t : a.sgd_b + (t - kNW);
            *wdst -= *a.lr_dev * acc;
        }
    }
    if (a.xc.world() > 1) {
        float s = 0.f;
        const bool ok = xchg::allreduce_small(a.xc, t, kNOut, acc, s);
        if (t < kNOut && ok) {
            const size_t idx = size_t(t < kNW ? a.w_off + t : a.b_off + (t - kNW));
            float* w = reinterpret_cast<float*>(a.xc.chan.region[a.xc.chan.rank] + a.xc.params_off) + idx;
            *w -= (*a.lr_dev / float(a.xc.world())) * s;
        }
    }
}

}  // namespace

bool tc_conv_tail_enabled() {
    // Opt-in (B200CNN_FUSED_C1=1).  Correct (the engine tests pass with it on, parity gate 6.1e-4), but measured SLOWER in
    // the step on B200: 76.0-77.5 us against 68.3-68.7 us.  The kernel itself takes 16.3 us from the end of its producer,
    // against 7.9 + 8.9 us for the two kernels it replaces — the boundary between them was already hidden by
    // programmatic dependent launch, and writing dz1 into the operand tile + a CTA-wide hand-over costs what the TMA
    // round trip did — while the second convolution's weight-gradient kernel on the side stream, which shares the SMs
    // with it, goes from 10.6 to 20 us (130 KB of shared memory and 384-512 threads per CTA on 128 SMs).
    static const bool on = [] {
        const char* v = getenv("B200CNN_FUSED_C1");
        return v && v[0] == '1';
    }();
    return on && current_backend() != B200CNN_BACKEND_SIMT && tc_conv1_bwd_weight_enabled();
}

bool tc_conv_tail_supported(const b200cnn_conv_shape* s2, const b200cnn_conv_shape* s1, int max_partials) {
    return tc_conv3x3s2_bwd_data_supported(s2) && s2->oh == 7 && s2->ow == 7 && s2->h == 14 && s2->w == 14 &&
           tc_conv1_bwd_weight_supported(s1, max_partials) && s1->n == s2->n;
}

// dz2: NHWC [n][7][7][32] or CHW [n][32][7][7]; a_prev: NHWC [n][14][14][16] or NULL; dx: NULL or NHWC [n][14][14][16];
// x: uint8 [n][28][28]; partials: ceil(n/2) * 160 floats; counter: one zero word.
int tc_conv2dx_conv1dw(const float* dz2, int dz_is_chw, const float* w2, const float* a_prev_nhwc, int act_prev, float* dx_nhwc,
                       const b200cnn_conv_shape* s2, const uint8_t* x, float* dw1, float* db1, float* partials, unsigned* counter,
                       const b200cnn_conv_shape* s1, float* sgd_w, float* sgd_b, const float* lr_dev, const TailXchg* xchg, long w_off,
                       long b_off, int early_w, cudaStream_t st) {
    if (!tc_conv_tail_supported(s2, s1, ceil_div(s1->n, kImgs))) return set_error(kInvalidValue, "tc_conv2dx_conv1dw: unsupported shapes");
    if (xchg && xchg->world() > 1 && (sgd_w || !lr_dev || !dw1 || !db1))
        return set_error(kInvalidValue, "tc_conv2dx_conv1dw: the in-kernel all-reduce needs dw/db in the gradient buffer and no local update");
    CUtensorMap tm;
    const uint64_t dims[4] = {uint64_t(OC), 7, 7, uint64_t(s2->n)};
    const uint64_t strides[4] = {0, uint64_t(OC) * 4, 7ull * OC * 4, 49ull * OC * 4};
    const uint32_t box[4] = {uint32_t(OC), 7, 7, uint32_t(kImgs)};
    // (any valid map when dz2 is CHW: the kernel does not touch it)
    if (int rc = encode_tmap_f32(&tm, dz_is_chw ? w2 : dz2, 4, dims, strides, box, nullptr, 128)) return rc;
    TailArgs a{dz_is_chw ? dz2 : nullptr, w2, a_prev_nhwc, dx_nhwc, x, partials, dw1, db1, counter, sgd_w, sgd_b, lr_dev, s2->n, act_prev,
               early_w, xchg ? *xchg : TailXchg{}, w_off, b_off};
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(tc_conv2dx_conv1dw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kSmem)));
        configured = true;
    }
    B200CNN_CUDA(launch_kernel(tc_conv2dx_conv1dw_kernel, dim3(ceil_div(s2->n, kImgs)), dim3(kThreads), kSmem, st, tm, a));
    B200CNN_CHECK_LAUNCH("tc_conv2dx_conv1dw");
    return 0;
}

}  // namespace b200cnn
