// 3x3 convolution on the tensor cores as an IMPLICIT GEMM fed by TMA — no im2col matrix exists in
// memory.  Activations between the convolutions are NHWC (channels innermost), so for one filter tap
// (ky,kx) the GEMM operand "all output pixels x input channels" is a strided 4-D box of the input:
//     dims (C, W, H, N), box (C, (OW-1)*s+1, (OH-1)*s+1, G) walked with traversal strides (1, s, s, 1)
// started at (0, kx-pad, ky-pad, n0).  TMA delivers it as [G*OH*OW rows][C floats] with out-of-bounds
// (padding) elements zero-filled, already in the swizzled K-major layout tcgen05.mma reads.  A 3x3
// convolution is then 9 taps x (C/8) MMAs accumulating into one TMEM tile.
//
//   forward        y[n][oc][oy][ox] = relu(b + sum_tap sum_c x[n][oy*s-p+ky][ox*s-p+kx][c] w[oc][c][ky][kx])
//                  A = input taps (TMA, K-major SW64), B = weights (gathered from OIHW into SW64 tiles),
//                  epilogue: bias + ReLU, stored CHW-flattened [n][oc*OH*OW + p] — the layout the
//                  reference's first FC layer indexes (/root/reference/cnn.c:116-121 over :179-206).
//   backward-data  dx[n][y][x][c] = relu'(a_prev) * sum_tap sum_oc dz[n][oy][ox][oc] w[oc][c][ky][kx]
//                  (stride 2: the four (y%2, x%2) parity classes are four small dense GEMMs over
//                  shifted views of dz; 4 TMA tiles, 9 tap-MMAs, 4 accumulators in TMEM.)
//
// Replaces Layer_feedForw_conv / the data half of Layer_feedBack_conv (cnn.c:175-247) and the
// reference's only kernel, conv_forward_kernel (CUDAcnn.cu:167-195: scalar fp64, batch 1, no ReLU).
#include <cstdlib>

#include "backend.cuh"
#include "common.cuh"
#include "tc_common.cuh"

namespace b200cnn {

namespace {

using namespace tc;

constexpr int kConvThreads = 192;  // warp 0: TMA, warp 1: TMEM + MMA, warps 2-5: weights gather + epilogue
constexpr int kTaps = 9;
constexpr int kCin = 16;           // input channels of the forward kernel: one 64-byte swizzle row per pixel
constexpr int kTileRows = 128;     // UMMA M

struct ConvFwdArgs {
    const float* w;     // OIHW
    const float* bias;  // [oc]
    float* y;           // [n][ldy], CHW-flattened per sample
    int ldy;
    int n, oc, oh, ow, pad, group;  // group = samples per CTA
    int act;
    int stride;
    int store_box;  // > 0: the epilogue stages the tile in shared memory and writes it with TMA boxes of this many floats
    int tiles;      // > 1: an image has more than 128 output pixels and is cut into `tiles` bands of oh_t output rows,
    int oh_t;       //      one band (and one sample) per CTA; tiles == 1: oh_t = oh and `group` samples per CTA
    int nhwc_out;   // output [n][oh][ow][oc] (feeds another NHWC layer) instead of CHW-flattened rows
    int early_w;    // the weights are not written by the kernel in front of this one: gather them BEFORE griddepcontrol.wait
};

// ---- forward -------------------------------------------------------------------------------------
template <int OC>
__global__ void __launch_bounds__(kConvThreads, 1)
tc_conv3x3_fwd_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_y, ConvFwdArgs a) {
    constexpr int kATile = kTileRows * kCin * 4;  // 8 KiB per tap (rows beyond group*oh*ow stay unused)
    constexpr int kBTile = OC * kCin * 4;         // OC rows x 64 B per tap
    constexpr int kTmemCols = OC < 32 ? 32 : OC;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_tiles = smem;
    uint8_t* b_tiles = smem + kTaps * kATile;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(b_tiles + kTaps * kBTile);
    uint64_t* tmem_full_bar = full_bar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    const int band = a.tiles > 1 ? int(blockIdx.x) % a.tiles : 0;      // which band of output rows of the image
    const int n0 = a.tiles > 1 ? int(blockIdx.x) / a.tiles : int(blockIdx.x) * a.group;
    const int pix = a.oh * a.ow;             // output pixels of an image
    const int tpix = a.oh_t * a.ow;          // ... of this CTA's band (= pix when tiles == 1)
    const int rows = min(a.group, a.n - n0) * tpix;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        if (a.store_box > 0) prefetch_tmap(&tmap_y);
        mbar_init(full_bar, 1);
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Weights: OIHW -> per-tap K-major [oc][c] tiles in the 64-byte-swizzle layout
    // (16-byte chunk index ^= (row >> 1) & 3, cute Swizzle<2,4,3>).  Every load of a thread is in flight at once: the
    // gather is a chain of L2 round trips otherwise.
    auto gather_weights = [&]() {
        constexpr int kW = OC * kCin * kTaps, kStride = kConvThreads - 64, kBatch = 12;
        for (int i0 = threadIdx.x - 64; i0 < kW; i0 += kBatch * kStride) {  // i walks OIHW linearly
            float v[kBatch];
#pragma unroll
            for (int u = 0; u < kBatch; u++) v[u] = (i0 + u * kStride < kW) ? __ldg(a.w + i0 + u * kStride) : 0.f;
#pragma unroll
            for (int u = 0; u < kBatch; u++) {
                const int i = i0 + u * kStride;
                if (i >= kW) break;
                const int t = i % kTaps, c = (i / kTaps) % kCin, oc = i / (kTaps * kCin);
                const int chunk = (c >> 2) ^ ((oc >> 1) & 3);
                *reinterpret_cast<float*>(b_tiles + t * kBTile + oc * 64 + chunk * 16 + (c & 3) * 4) = v[u];
            }
        }
        fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor-core (async) proxy
    };
    if (a.early_w && warp >= 2) gather_weights();  // overlaps the tail of the kernel in front
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0) {
            // One barrier for the whole K extent: 9 tap boxes of group*oh*ow pixels x 64 B.
            mbar_arrive_expect_tx(full_bar, uint32_t(kTaps * a.group * tpix * kCin * 4));
#pragma unroll
            for (int t = 0; t < kTaps; t++)
                tma_load_4d(a_tiles + t * kATile, &tmap_x, full_bar, 0, t % 3 - a.pad, band * a.oh_t * a.stride + t / 3 - a.pad, n0);
        }
    } else if (warp >= 2 && !a.early_w) {
        gather_weights();
    }
    __syncthreads();

    if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(kTileRows, OC, false, false);
            mbar_wait(full_bar, 0);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_tiles), b_addr = smem_u32(b_tiles);
#pragma unroll
            for (int t = 0; t < kTaps; t++)
#pragma unroll
                for (int kk = 0; kk < kCin / 8; kk++) {
                    const uint64_t da = make_smem_desc(a_addr + t * kATile + kk * 32, 16, 512, kSw64);
                    const uint64_t db = make_smem_desc(b_addr + t * kBTile + kk * 32, 16, 512, kSw64);
                    umma_tf32(tmem_base, da, db, idesc, (t | kk) != 0);
                }
            umma_commit(tmem_full_bar);
        }
    } else if (warp >= 2) {
        const int q = warp & 3;
        const int r = 32 * q + lane;
        float bias[OC];  // fetched before the accumulator is ready
#pragma unroll
        for (int i = 0; i < OC; i++) bias[i] = __ldg(a.bias + i);
        mbar_wait(tmem_full_bar, 0);
        tcgen05_fence_after();
        const bool live = r < rows;
        const int g = r / tpix, p = band * tpix + r % tpix;  // sample of the group, pixel of the image
        if (a.nhwc_out) {
            // channels innermost: a thread owns all OC values of its pixel -> OC / 4 16-byte stores
            float* out = a.y + (size_t(n0 + g) * pix + p) * OC;
#pragma unroll
            for (int c0 = 0; c0 < OC; c0 += 16) {
                float v[16];
                __syncwarp();
                tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(c0), v);
                if (live) {
#pragma unroll
                    for (int i = 0; i < 16; i += 4)
                        *reinterpret_cast<float4*>(out + c0 + i) =
                            make_float4(apply_act(a.act, v[i] + bias[c0 + i]), apply_act(a.act, v[i + 1] + bias[c0 + i + 1]),
                                        apply_act(a.act, v[i + 2] + bias[c0 + i + 2]), apply_act(a.act, v[i + 3] + bias[c0 + i + 3]));
                }
            }
        } else if (a.store_box > 0) {
            // Epilogue through shared memory + TMA store: the A tiles are dead once the accumulator is complete, so
            // the CHW-flattened output tile [group][oc * pix] is assembled there as boxes of `store_box` floats per
            // sample ([box][sample][store_box], the dense layout a 2-D TMA box reads) and leaves the SM as a few bulk
            // tensor stores instead of 32 strided 4-byte stores per thread.
            float* stage = reinterpret_cast<float*>(a_tiles);
            const int bw = a.store_box, per_box = a.group * bw;
#pragma unroll
            for (int c0 = 0; c0 < OC; c0 += 32) {
                float v[32];
                __syncwarp();
                if (OC - c0 >= 32) {
                    tmem_ld_32x32b_x32(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(c0), v);
                } else {
                    float h[16];
                    tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(c0), h);
#pragma unroll
                    for (int i = 0; i < 16; i++) v[i] = h[i];
                }
                if (live) {
#pragma unroll
                    for (int i = 0; i < (OC - c0 >= 32 ? 32 : 16); i++) {
                        const int e = (c0 + i) * pix + p;  // position inside the sample's CHW-flattened output
                        const int b = e / bw;
                        stage[b * per_box + g * bw + (e - b * bw)] = apply_act(a.act, v[i] + bias[c0 + i]);
                    }
                }
            }
            fence_proxy_async_smem();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (threadIdx.x == 64) {
                const int nbox = a.oc * pix / bw;
                for (int b = 0; b < nbox; b++) tma_store_2d(&tmap_y, stage + b * per_box, b * bw, n0);  // rows past n are clipped
                tma_store_commit_and_wait();
            }
        } else {
            float* out = a.y + size_t(n0 + g) * a.ldy + p;
#pragma unroll
            for (int c0 = 0; c0 < OC; c0 += 16) {
                float v[16];
                __syncwarp();
                tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(c0), v);
                if (live) {
#pragma unroll
                    for (int i = 0; i < 16; i++)
                        out[size_t(c0 + i) * pix] = apply_act(a.act, v[i] + bias[c0 + i]);
                }
            }
        }
        __syncwarp();
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

// ---- backward-data (3x3, stride 2, pad 1) ------------------------------------------------------------
struct ConvBwdArgs {
    const float* dz_chw;  // if non-NULL: dz as [n][oc][oh][ow] (what the FC layer above produces); gathered by threads
    const float* w;       // OIHW [oc][c][3][3]
    const float* a_prev;  // NHWC [n][h][w][c] activation that fed the conv (ReLU mask), or NULL
    float* dx;            // NHWC [n][h][w][c]
    int n, oh, ow, h, w_in, group;
    int act_prev;
    int early_w;  // as in ConvFwdArgs
};

// Output-gradient channels OC = 32 (one 128-byte swizzle row per pixel), input channels C = 16.
template <int OC, int C>
__global__ void __launch_bounds__(kConvThreads, 1)
tc_conv3x3s2_bwd_data_kernel(const __grid_constant__ CUtensorMap tmap_dz, ConvBwdArgs a) {
    static_assert(OC == 32, "one 128-byte row of dz per output pixel");
    constexpr int kATile = kTileRows * OC * 4;  // 16 KiB per shift (dj, di)
    constexpr int kBTile = C * OC * 4;          // [c rows][oc] per tap, 128-byte rows
    constexpr int kTmemCols = 64 < 4 * C ? 4 * C : 64;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_tiles = smem;                  // 4 shifted views of dz
    uint8_t* b_tiles = smem + 4 * kATile;     // 9 taps
    float* stage = reinterpret_cast<float*>(b_tiles + kTaps * kBTile);  // [rows][OC] staging of CHW gradients
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(stage) + kTileRows * OC * 4);
    uint64_t* tmem_full_bar = full_bar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    const int n0 = blockIdx.x * a.group;
    const int pix = a.oh * a.ow;
    const int rows = min(a.group, a.n - n0) * pix;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_dz);
        mbar_init(full_bar, 1);
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Weights per tap as the B operand [n = c][k = oc], K-major 128-byte rows, 128B swizzle (16-byte chunk ^= row & 7);
    // every load of a thread in flight at once.
    auto gather_weights = [&]() {
        constexpr int kW = OC * C * kTaps, kStride = kConvThreads - 64, kBatch = 12;
        for (int i0 = threadIdx.x - 64; i0 < kW; i0 += kBatch * kStride) {
            float v[kBatch];
#pragma unroll
            for (int u = 0; u < kBatch; u++) v[u] = (i0 + u * kStride < kW) ? __ldg(a.w + i0 + u * kStride) : 0.f;
#pragma unroll
            for (int u = 0; u < kBatch; u++) {
                const int i = i0 + u * kStride;
                if (i >= kW) break;
                const int t = i % kTaps, c = (i / kTaps) % C, oc = i / (kTaps * C);
                const int chunk = (oc >> 2) ^ (c & 7);
                *reinterpret_cast<float*>(b_tiles + t * kBTile + c * 128 + chunk * 16 + (oc & 3) * 4) = v[u];
            }
        }
    };
    if (a.early_w && warp >= 2) gather_weights();  // overlaps the tail of the kernel in front
    pdl_wait();
    pdl_launch_dependents();

    if (a.dz_chw) {
        // dz arrives CHW (the FC layer's gradient w.r.t. its flattened input): transpose it on the fly into
        // the four shifted K-major tiles, 128-byte rows, 128B swizzle.  Reads are coalesced along the pixel
        // index; each thread writes one 16-byte chunk (4 channels of one pixel), conflict-free thanks to
        // the swizzle.
        // Stage 1: the group's samples are contiguous in memory ([n][OC][pix]); one round of coalesced
        // float4 loads brings them into shared memory.
        const int nfl = rows * OC;  // floats of this group
        {
            const float4* src = reinterpret_cast<const float4*>(a.dz_chw + size_t(n0) * OC * pix);
            for (int i = threadIdx.x; i < nfl / 4; i += kConvThreads) reinterpret_cast<float4*>(stage)[i] = __ldg(src + i);
            for (int i = (nfl & ~3) + threadIdx.x; i < nfl; i += kConvThreads) stage[i] = __ldg(a.dz_chw + size_t(n0) * OC * pix + i);
        }
        __syncthreads();
        // Stage 2: shared -> shared transpose into the four shifted, swizzled K-major tiles.
        // Thread r < 128 owns tile row r (one output pixel): its (sample, y, x) is computed once, the loop over
        // (shift, channel group) has no divisions.  Threads 128..191 take the second half of the channel groups.
        {
            const int r = threadIdx.x % kTileRows, half = threadIdx.x / kTileRows;  // half in {0, 1}
            if (r < rows) {
                const int g = r / pix, p = r % pix, j = p / a.ow, i = p % a.ow;
                const float* base = stage + g * OC * pix;
                const bool two = kConvThreads >= 2 * kTileRows;  // false: 192 threads => rows 0..63 get a helper
                (void)two;
                for (int sft = 0; sft < 4; sft++) {
                    const int jj = j + (sft >> 1), ix = i + (sft & 1);
                    const bool inb = jj < a.oh && ix < a.ow;
                    const float* src = base + jj * a.ow + ix;
                    uint8_t* dst = a_tiles + sft * kATile + r * 128;
                    // rows 0..63 have two threads (r and r + 128): split the 8 channel groups between them
                    const int lo = (r < kConvThreads - kTileRows) ? half * (OC / 8) : 0;
                    const int hi = (r < kConvThreads - kTileRows) ? lo + OC / 8 : (half == 0 ? OC / 4 : 0);
                    for (int oc4 = lo; oc4 < hi; oc4++) {
                        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (inb) {
                            const float* s4 = src + oc4 * 4 * pix;
                            v = make_float4(s4[0], s4[pix], s4[2 * pix], s4[3 * pix]);
                        }
                        *reinterpret_cast<float4*>(dst + ((oc4 ^ (r & 7)) << 4)) = v;
                    }
                }
            }
        }
    } else if (warp == 0) {
        if (lane == 0) {
            mbar_arrive_expect_tx(full_bar, uint32_t(4 * a.group * pix * OC * 4));
#pragma unroll
            for (int sft = 0; sft < 4; sft++)  // shift (dj, di): rows (j, i) read dz[n][j+dj][i+di][:], zero past the edge
                tma_load_4d(a_tiles + sft * kATile, &tmap_dz, full_bar, 0, sft & 1, sft >> 1, n0);
        }
    }
    if (warp >= 2 && !a.early_w) gather_weights();
    fence_proxy_async_smem();
    __syncthreads();

    if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(kTileRows, C, false, false);
            if (!a.dz_chw) mbar_wait(full_bar, 0);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_tiles), b_addr = smem_u32(b_tiles);
            // Input pixel (2j+py, 2i+px) receives tap (ky,kx) from output pixel (j+dj, i+di):
            //   py = 0: ky = 1, dj = 0          py = 1: ky = 0, dj = 1  and  ky = 2, dj = 0     (same for x)
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
                        const int sft = dj * 2 + di, t = ky * 3 + kx;
#pragma unroll
                        for (int kk = 0; kk < OC / 8; kk++) {
                            const uint64_t da = make_smem_desc(a_addr + sft * kATile + kk * 32, 16, 1024, kSw128);
                            const uint64_t db = make_smem_desc(b_addr + t * kBTile + kk * 32, 16, 1024, kSw128);
                            umma_tf32(tmem_base + cls * C, da, db, idesc, !(first && kk == 0));
                        }
                        first = false;
                    }
                }
            }
            umma_commit(tmem_full_bar);
        }
    } else if (warp >= 2) {
        const int q = warp & 3;
        const int r = 32 * q + lane;
        const bool live = r < rows;
        const int g = r / pix, p = r % pix, j = p / a.ow, i = p % a.ow;
        // Prefetch the ReLU mask source (the previous layer's activations at this thread's four input pixels) before
        // waiting for the accumulators: 16 independent loads in flight instead of four exposed round trips.
        float4 msk[4][C / 4];
        if (live && a.a_prev) {
#pragma unroll
            for (int cls = 0; cls < 4; cls++) {
                const int y = 2 * j + (cls >> 1), x = 2 * i + (cls & 1);
                const float4* ap = reinterpret_cast<const float4*>(a.a_prev + ((size_t(n0 + g) * a.h + y) * a.w_in + x) * C);
#pragma unroll
                for (int k4 = 0; k4 < C / 4; k4++) msk[cls][k4] = (y < a.h && x < a.w_in) ? __ldg(ap + k4) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        mbar_wait(tmem_full_bar, 0);
        tcgen05_fence_after();
#pragma unroll
        for (int cls = 0; cls < 4; cls++) {
            float v[16];
            __syncwarp();
            tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(cls * C), v);
            const int y = 2 * j + (cls >> 1), x = 2 * i + (cls & 1);
            if (live && y < a.h && x < a.w_in) {
                const size_t off = ((size_t(n0 + g) * a.h + y) * a.w_in + x) * C;
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
                float4* dp = reinterpret_cast<float4*>(a.dx + off);
#pragma unroll
                for (int k4 = 0; k4 < C / 4; k4++) dp[k4] = make_float4(v[4 * k4], v[4 * k4 + 1], v[4 * k4 + 2], v[4 * k4 + 3]);
            }
        }
        __syncwarp();
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

// ---- backward-data (3x3, stride 1, pad 1): the pooled topology's second convolution ------------------------------
// dx[n][y][x][c] = act'(a_prev) * sum_tap sum_oc dz[n][y+1-ky][x+1-kx][oc] w[oc][c][ky][kx]: a plain 3x3 convolution of
// dz with the taps mirrored.  Same implicit GEMM as the forward kernel — per tap ONE TMA box of the NHWC gradient
// (32 channels = one 128-byte swizzle row per pixel, out-of-bounds = zero padding), B = the tap's [c][oc] weights —
// 9 taps x 4 MMAs into one accumulator.  An image has 196 pixels, so it is cut into bands of whole rows (<= 128 pixels).
struct ConvS1BwdArgs {
    const float* w;       // OIHW [oc][c][3][3]
    const float* a_prev;  // NHWC [n][h][w][c] or NULL
    float* dx;            // NHWC [n][h][w][c]
    int n, h, w_in, tiles, h_t;
    int act_prev;
    int early_w;  // as in ConvFwdArgs
};

template <int OC, int C>
__global__ void __launch_bounds__(kConvThreads, 1)
tc_conv3x3s1_bwd_data_kernel(const __grid_constant__ CUtensorMap tmap_dz, ConvS1BwdArgs a) {
    static_assert(OC == 32, "one 128-byte row of dz per pixel");
    constexpr int kATile = kTileRows * OC * 4;  // 16 KiB per tap
    constexpr int kBTile = C * OC * 4;          // [c rows][oc] per tap, 128-byte rows
    constexpr int kTmemCols = 32;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_tiles = smem;
    uint8_t* b_tiles = smem + kTaps * kATile;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(b_tiles + kTaps * kBTile);
    uint64_t* tmem_full_bar = full_bar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    const int band = int(blockIdx.x) % a.tiles, n0 = int(blockIdx.x) / a.tiles;
    const int rows = a.h_t * a.w_in;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_dz);
        mbar_init(full_bar, 1);
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Weights per tap as the B operand [n = c][k = oc], K-major 128-byte rows, 128B swizzle (16-byte chunk ^= row & 7);
    // every load of a thread in flight at once.
    auto gather_weights = [&]() {
        constexpr int kW = OC * C * kTaps, kStride = kConvThreads - 64, kBatch = 12;
        for (int i0 = threadIdx.x - 64; i0 < kW; i0 += kBatch * kStride) {
            float v[kBatch];
#pragma unroll
            for (int u = 0; u < kBatch; u++) v[u] = (i0 + u * kStride < kW) ? __ldg(a.w + i0 + u * kStride) : 0.f;
#pragma unroll
            for (int u = 0; u < kBatch; u++) {
                const int i = i0 + u * kStride;
                if (i >= kW) break;
                const int t = i % kTaps, c = (i / kTaps) % C, oc = i / (kTaps * C);
                const int chunk = (oc >> 2) ^ (c & 7);
                *reinterpret_cast<float*>(b_tiles + t * kBTile + c * 128 + chunk * 16 + (oc & 3) * 4) = v[u];
            }
        }
    };
    if (a.early_w && warp >= 2) {  // overlaps the tail of the kernel in front
        gather_weights();
        fence_proxy_async_smem();
    }
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0) {
            mbar_arrive_expect_tx(full_bar, uint32_t(kTaps * rows * OC * 4));
#pragma unroll
            for (int t = 0; t < kTaps; t++)  // tap (ky, kx) reads dz at (y + 1 - ky, x + 1 - kx)
                tma_load_4d(a_tiles + t * kATile, &tmap_dz, full_bar, 0, 1 - t % 3, band * a.h_t + 1 - t / 3, n0);
        }
    } else if (warp >= 2 && !a.early_w) {
        gather_weights();
        fence_proxy_async_smem();
    }
    __syncthreads();

    if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(kTileRows, C, false, false);
            mbar_wait(full_bar, 0);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_tiles), b_addr = smem_u32(b_tiles);
#pragma unroll
            for (int t = 0; t < kTaps; t++)
#pragma unroll
                for (int kk = 0; kk < OC / 8; kk++)
                    umma_tf32(tmem_base, make_smem_desc(a_addr + t * kATile + kk * 32, 16, 1024, kSw128),
                              make_smem_desc(b_addr + t * kBTile + kk * 32, 16, 1024, kSw128), idesc, (t | kk) != 0);
            umma_commit(tmem_full_bar);
        }
    } else if (warp >= 2) {
        const int q = warp & 3;
        const int r = 32 * q + lane;
        const bool live = r < rows;
        const size_t off = (size_t(n0) * a.h * a.w_in + size_t(band) * rows + r) * C;
        float4 msk[C / 4];
        if (live && a.a_prev) {
#pragma unroll
            for (int k4 = 0; k4 < C / 4; k4++) msk[k4] = __ldg(reinterpret_cast<const float4*>(a.a_prev + off) + k4);
        }
        mbar_wait(tmem_full_bar, 0);
        tcgen05_fence_after();
        float v[16];
        tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16), v);
        if (live) {
            if (a.a_prev) {
#pragma unroll
                for (int k4 = 0; k4 < C / 4; k4++) {
                    v[4 * k4 + 0] *= act_grad_from_output(a.act_prev, msk[k4].x);
                    v[4 * k4 + 1] *= act_grad_from_output(a.act_prev, msk[k4].y);
                    v[4 * k4 + 2] *= act_grad_from_output(a.act_prev, msk[k4].z);
                    v[4 * k4 + 3] *= act_grad_from_output(a.act_prev, msk[k4].w);
                }
            }
            float4* dp = reinterpret_cast<float4*>(a.dx + off);
#pragma unroll
            for (int k4 = 0; k4 < C / 4; k4++) dp[k4] = make_float4(v[4 * k4], v[4 * k4 + 1], v[4 * k4 + 2], v[4 * k4 + 3]);
        }
        __syncwarp();
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

}  // namespace

// Bands an image is cut into so that a band has at most 128 output pixels (0: no such cut with whole rows).
static int fwd_tiles(const b200cnn_conv_shape* s) {
    if (s->oh * s->ow <= kTileRows) return 1;
    for (int t = 2; t <= s->oh; t++)
        if (s->oh % t == 0 && (s->oh / t) * s->ow <= kTileRows) return t;
    return 0;
}

bool tc_conv3x3_fwd_supported(const b200cnn_conv_shape* s) {
    return s->k == 3 && s->c == kCin && (s->oc == 16 || s->oc == 32 || s->oc == 64) && fwd_tiles(s) > 0 && s->stride >= 1 &&
           s->stride <= 8 && (s->ow - 1) * s->stride + 1 <= 256 && (s->oh - 1) * s->stride + 1 <= 256;
}

// x: NHWC [n][h][w][16]; y: [n][ldy] with the sample's outputs CHW-flattened (ldy >= oc*oh*ow).
int tc_conv3x3_fwd(const float* x_nhwc, const float* w, const float* b, float* y, int ldy, const b200cnn_conv_shape* s,
                   int act, cudaStream_t st, int nhwc_out, int early_w) {
    if (!tc_conv3x3_fwd_supported(s)) return set_error(kInvalidValue, "tc_conv3x3_fwd: unsupported shape");
    if (tc_conv3x3s1_halo_fwd_supported(s))  // stride 1: one operand load with a halo instead of nine boxes (tc_conv_halo.cu)
        return tc_conv3x3s1_halo_fwd(x_nhwc, w, b, y, ldy, s, act, st, nhwc_out, early_w);
    const int pix = s->oh * s->ow;
    const int tiles = fwd_tiles(s), oh_t = s->oh / tiles;
    const int group = tiles > 1 ? 1 : kTileRows / pix;
    CUtensorMap tm;
    const uint64_t dims[4] = {uint64_t(s->c), uint64_t(s->w), uint64_t(s->h), uint64_t(s->n)};
    const uint64_t strides[4] = {0, uint64_t(s->c) * 4, uint64_t(s->w) * s->c * 4, uint64_t(s->h) * s->w * s->c * 4};
    const uint32_t box[4] = {uint32_t(s->c), uint32_t((s->ow - 1) * s->stride + 1), uint32_t((oh_t - 1) * s->stride + 1),
                             uint32_t(group)};
    const uint32_t estr[4] = {1, uint32_t(s->stride), uint32_t(s->stride), 1};
    if (int rc = encode_tmap_f32(&tm, x_nhwc, 4, dims, strides, box, estr, 64)) return rc;
    // Output through TMA: y seen as [n][oc * pix] with row stride ldy, written in boxes of `bw` floats x group samples
    // (bw = the largest divisor of oc * pix that is a multiple of 4 and <= 256).  Needs 16-byte aligned rows.
    int bw = 0;
    const int per_sample = s->oc * pix;
    static const bool tma_store = [] {
        // Opt-in (B200CNN_TMA_STORE=1).  Correct (the whole GPU suite passes with it on) but measured SLOWER in the
        // step on B200: 73.8 vs 71.4 us.  The plain stores are fire-and-forget and coalesce (consecutive lanes =
        // consecutive pixels of one channel); the TMA path adds a shared-memory pass, a proxy fence, a CTA barrier and
        // one thread issuing seven bulk stores at ~100 ns each to the tail of a kernel whose whole life is ~7 us.
        const char* v = getenv("B200CNN_TMA_STORE");
        return v && v[0] == '1';
    }();
    if (tma_store && !nhwc_out && tiles == 1 && ldy % 4 == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0 && size_t(group) * per_sample * 4 <= size_t(kTaps) * kTileRows * kCin * 4)
        for (int c = 256; c >= 4; c -= 4)
            if (per_sample % c == 0) {
                bw = c;
                break;
            }
    CUtensorMap tmy = tm;
    if (bw > 0) {
        const uint64_t ydims[2] = {uint64_t(per_sample), uint64_t(s->n)}, ystr[2] = {0, uint64_t(ldy) * 4};
        const uint32_t ybox[2] = {uint32_t(bw), uint32_t(group)};
        if (int rc = encode_tmap_f32(&tmy, y, 2, ydims, ystr, ybox, nullptr, 0)) return rc;
    }
    ConvFwdArgs a{w, b, y, ldy, s->n, s->oc, s->oh, s->ow, s->pad, group, act, s->stride, bw, tiles, oh_t, nhwc_out, early_w};
    const int grid = tiles > 1 ? s->n * tiles : ceil_div(s->n, group);
#define B200CNN_LAUNCH_FWD(OCV)                                                                                   \
    do {                                                                                                          \
        auto kern = tc_conv3x3_fwd_kernel<OCV>;                                                                   \
        constexpr size_t smem = size_t(kTaps) * (kTileRows * kCin * 4 + OCV * kCin * 4) + 1024 + 64;               \
        static bool configured = false;                                                                           \
        if (!configured) {                                                                                        \
            B200CNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));      \
            configured = true;                                                                                    \
        }                                                                                                         \
        B200CNN_CUDA(launch_kernel(kern, dim3(grid), dim3(kConvThreads), smem, st, tm, tmy, a));                     \
    } while (0)
    if (s->oc == 16) B200CNN_LAUNCH_FWD(16);
    else if (s->oc == 32) B200CNN_LAUNCH_FWD(32);
    else B200CNN_LAUNCH_FWD(64);
#undef B200CNN_LAUNCH_FWD
    B200CNN_CHECK_LAUNCH("tc_conv3x3_fwd");
    return 0;
}

bool tc_conv3x3s2_bwd_data_supported(const b200cnn_conv_shape* s) {
    return s->k == 3 && s->stride == 2 && s->pad == 1 && s->c == 16 && s->oc == 32 && s->oh * s->ow <= kTileRows &&
           s->h == 2 * s->oh && s->w == 2 * s->ow;
}

// dz: NHWC [n][oh][ow][32], or CHW [n][32][oh][ow] when dz_is_chw; a_prev / dx: NHWC [n][h][w][16].
int tc_conv3x3s2_bwd_data(const float* dz, int dz_is_chw, const float* w, const float* a_prev_nhwc, int act_prev,
                          float* dx_nhwc, const b200cnn_conv_shape* s, cudaStream_t st, int early_w) {
    const float* dz_nhwc = dz;
    if (!tc_conv3x3s2_bwd_data_supported(s)) return set_error(kInvalidValue, "tc_conv3x3s2_bwd_data: unsupported shape");
    const int pix = s->oh * s->ow;
    const int group = kTileRows / pix;
    CUtensorMap tm;
    const uint64_t dims[4] = {uint64_t(s->oc), uint64_t(s->ow), uint64_t(s->oh), uint64_t(s->n)};
    const uint64_t strides[4] = {0, uint64_t(s->oc) * 4, uint64_t(s->ow) * s->oc * 4, uint64_t(pix) * s->oc * 4};
    const uint32_t box[4] = {uint32_t(s->oc), uint32_t(s->ow), uint32_t(s->oh), uint32_t(group)};
    if (int rc = encode_tmap_f32(&tm, dz_nhwc, 4, dims, strides, box, nullptr, 128)) return rc;
    ConvBwdArgs a{dz_is_chw ? dz : nullptr, w, a_prev_nhwc, dx_nhwc, s->n, s->oh, s->ow, s->h, s->w, group, act_prev, early_w};
    auto kern = tc_conv3x3s2_bwd_data_kernel<32, 16>;
    constexpr size_t smem = 4 * size_t(kTileRows) * 32 * 4 + kTaps * 16 * 32 * 4 + size_t(kTileRows) * 32 * 4 + 1024 + 64;
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        configured = true;
    }
    B200CNN_CUDA(launch_kernel(kern, dim3(ceil_div(s->n, group)), dim3(kConvThreads), smem, st, tm, a));
    B200CNN_CHECK_LAUNCH("tc_conv3x3s2_bwd_data");
    return 0;
}

static int s1_tiles(const b200cnn_conv_shape* s) {
    for (int t = 1; t <= s->h; t++)
        if (s->h % t == 0 && (s->h / t) * s->w <= kTileRows) return t;
    return 0;
}

bool tc_conv3x3s1_bwd_data_supported(const b200cnn_conv_shape* s) {
    return s->k == 3 && s->stride == 1 && s->pad == 1 && s->c == 16 && s->oc == 32 && s->oh == s->h && s->ow == s->w && s->w <= 128 &&
           s1_tiles(s) > 0;
}

// dz: NHWC [n][h][w][32]; a_prev / dx: NHWC [n][h][w][16].
int tc_conv3x3s1_bwd_data(const float* dz_nhwc, const float* w, const float* a_prev_nhwc, int act_prev, float* dx_nhwc,
                          const b200cnn_conv_shape* s, cudaStream_t st, int early_w) {
    if (!tc_conv3x3s1_bwd_data_supported(s)) return set_error(kInvalidValue, "tc_conv3x3s1_bwd_data: unsupported shape");
    if (tc_conv3x3s1_halo_bwd_data_supported(s)) return tc_conv3x3s1_halo_bwd_data(dz_nhwc, w, a_prev_nhwc, act_prev, dx_nhwc, s, st, early_w);
    const int tiles = s1_tiles(s), h_t = s->h / tiles;
    CUtensorMap tm;
    const uint64_t dims[4] = {uint64_t(s->oc), uint64_t(s->ow), uint64_t(s->oh), uint64_t(s->n)};
    const uint64_t strides[4] = {0, uint64_t(s->oc) * 4, uint64_t(s->ow) * s->oc * 4, uint64_t(s->oh) * s->ow * s->oc * 4};
    const uint32_t box[4] = {uint32_t(s->oc), uint32_t(s->ow), uint32_t(h_t), 1};
    if (int rc = encode_tmap_f32(&tm, dz_nhwc, 4, dims, strides, box, nullptr, 128)) return rc;
    ConvS1BwdArgs a{w, a_prev_nhwc, dx_nhwc, s->n, s->h, s->w, tiles, h_t, act_prev, early_w};
    auto kern = tc_conv3x3s1_bwd_data_kernel<32, 16>;
    constexpr size_t smem = size_t(kTaps) * (kTileRows * 32 * 4 + 16 * 32 * 4) + 1024 + 64;
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        configured = true;
    }
    B200CNN_CUDA(launch_kernel(kern, dim3(s->n * tiles), dim3(kConvThreads), smem, st, tm, a));
    B200CNN_CHECK_LAUNCH("tc_conv3x3s1_bwd_data");
    return 0;
}

}  // namespace b200cnn
