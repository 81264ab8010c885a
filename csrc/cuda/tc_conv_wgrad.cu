// Weight gradient of the 3x3 / stride-2 / pad-1 convolution (16 -> 32 channels, 14x14 -> 7x7) on the tensor
// cores, as a split-K implicit GEMM:
//     dW[oc][c][ky][kx] = sum over (n, oy, ox) of dz[n][oy][ox][oc] * x[n][2oy-1+ky][2ox-1+kx][c]
// i.e. D[M = (kx, c)][N = oc] = sum_K A[K][M] * B[K][N] with K = output pixels, for each ky.  Both operands are
// "MN-major" (the K index is the ROW of the staged tile), which is exactly what TMA delivers without any im2col:
//   A  x is NHWC with 16 channels, so two horizontally adjacent pixels are 32 contiguous floats = one 128-byte
//      row.  Seen as dims (32 floats, W/2 pairs, H, N) a box (32, 7, 13 walked with stride 2, 4 images) started
//      at pair 0 / row ky-1 is [196 pixels][taps kx=1 | kx=2], and started at pair -1 it is [196 pixels]
//      [unused | tap kx=0]; out-of-bounds pairs / rows are the zero padding.  Two boxes = one 64-row M extent.
//   B  dz [196 pixels][32 oc]: one TMA box when dz is NHWC; when it arrives CHW (what the FC layer above
//      writes) the epilogue warps repack it: coalesced float4 loads of one image into a 6 KB buffer, then a
//      conflict-free shared -> shared transpose into the swizzled tile, the next image's loads already in flight.
// A CTA owns 8 images and runs them as two phases of 4 (196 K rows + 4 zero rows = 25 tcgen05.mma of M=64, N=32,
// K=8 each) through ONE set of tiles, so it needs 83 KB of shared memory and fits on an SM next to the CTAs of the
// data-gradient kernel and of the first convolution's weight-gradient kernel that run at the same time.
// Grid (3 ky, 1, n/8): every CTA stores its 64x32 partial tile, and a second small kernel (chained with
// programmatic dependent launch) adds the n/8 partials in index order and scatters dW / db in the reference's
// OIHW layout: fixed summation order, run-to-run identical results.  (Measured alternative, kept out: reducing
// inside the kernel through DSMEM in clusters of 8 + a "last leader" pass cost 13 us of barriers, fences and
// one-SM funnelling against ~3 us for the second kernel.)
// Bias gradient: column sums of dz, carried in row 0 of the partial tile (an unused tap row).
//
// Stride 1 (kMode 1: the pooled topology's second convolution, 14x14 -> 14x14): the same kernel over other views.
// Output pixel x reads input pixels x-1, x, x+1, so the pair trick needs the outputs split by the parity of x:
//   x = 2i:     [kx0 | kx1'] = odd-aligned pair i-1 = pixels (2i-1, 2i),   [kx1 | kx2] = even-aligned pair i
//   x = 2i + 1: [kx0 | kx1]  = even-aligned pair i,                        [kx1' | kx2] = odd-aligned pair i = (2i+1, 2i+2)
// (kx1' is a duplicate that the reduction ignores).  "Odd-aligned pairs" are a second tensor map over the same memory
// shifted by one pixel, with 6 pairs per row: pair -1 and pair 6 are out of bounds = the zero padding.  dz is read
// through a 5-D map (32 oc, parity, 7, 14, n), so its rows come in the same (y, i) order.  A phase is 2 images
// (2 x 98 = 196 K rows, again), a CTA owns 4 images of one parity: grid (3 ky, 2 parities, n/4).
//
// Replaces the weight half of Layer_feedBack_conv (/root/reference/cnn.c:236,242) for the second convolution.
#include <cstdlib>

#include "backend.cuh"
#include "common.cuh"
#include "tc_common.cuh"

namespace b200cnn {

namespace {

using namespace tc;

constexpr int kWgThreads = 192;               // warp 0: TMA, warp 1: TMEM + MMA, warps 2-5: dz repack + epilogue
constexpr int kIpc = 8;                       // images per CTA ...
constexpr int kPhaseImgs = 4;                 // ... in two phases of 4
constexpr int kPix = 49;                      // 7 x 7 output pixels
constexpr int kBoxRows = kPhaseImgs * kPix;   // 196 K rows written per phase
constexpr int kSteps = (kBoxRows + 7) / 8;    // 25 MMA steps (rows 196..199 are zero)
constexpr int kTileBytes = kSteps * 8 * 128;  // 25,600: one [200][32 floats] tile
constexpr int kBoxBytes = kBoxRows * 128;
constexpr int kImgFloats = 32 * kPix;         // one image of dz
constexpr int kPartFloats = 64 * 32;          // partial tile: rows (kx+1)*16 + c, row 0 = bias gradient
constexpr int kOC = 32, kC = 16;
constexpr size_t kWgSmem = 3 * size_t(kTileBytes) + kImgFloats * 4 + 4 * 32 * 4 + 64 + 1024;

struct WgradArgs {
    const float* dz_chw;   // non-NULL: dz as [n][32][49]
    const float* dz_nhwc;  // non-NULL: dz as [n][49][32] (also behind tmap_dz)
    float* part;           // [3][n/8][64][32]
    float* dw;             // OIHW [32][16][3][3]
    float* db;             // [32]
    int n;
    long long* dbg;        // instrumentation (B200CNN_WG_DEBUG=1): clock64 stamps of CTAs (ky, z < 4), [ky*4+z][16]
};

__device__ __forceinline__ void tma_load_5d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}

#define WG_STAMP(i) do { if (a.dbg && blockIdx.z < 4) a.dbg[(ky * 4 + blockIdx.z) * 16 + (i)] = clock64(); } while (0)
__device__ __forceinline__ void epilogue_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// kMode 0: stride 2 (tmap_x2 unused); kMode 1: stride 1 (tmap_x = even-aligned pairs, tmap_x2 = odd-aligned pairs).
// (Registers capped at 85: this kernel shares the SMs with the kernels of the critical path — two of its CTAs at the
//  146 registers the compiler would take leave no room for a CTA of the first convolution's weight-gradient kernel.)
template <int kMode>
__global__ void __launch_bounds__(kWgThreads, 4)
tc_conv3x3_bwd_weight_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_x2,
                             const __grid_constant__ CUtensorMap tmap_dz, WgradArgs a) {
    constexpr int kPhaseImgs = kMode ? 2 : b200cnn::kPhaseImgs;  // images per phase (196 K rows either way)
    constexpr int kIpc = 2 * kPhaseImgs;                         // images per CTA
    constexpr int kRowsPerImg = kMode ? 196 : kPix;              // dz rows of an image (all parities)
    constexpr int kBiasRow = kMode ? 16 : 0;                     // a tile row the reduction does not read as a weight
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_tile = smem;                        // 2 groups: [unused | kx0], [kx1 | kx2]
    uint8_t* b_tile = smem + 2 * kTileBytes;       // dz
    float* raw = reinterpret_cast<float*>(smem + 3 * kTileBytes);  // one CHW image of dz
    float* dbp = raw + kImgFloats;                 // [4 warps][32] bias-gradient partials
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(dbp + 4 * 32);  // TMA boxes of a phase have landed
    uint64_t* b_bar = full_bar + 1;                // repacked dz of a phase is in place (4 warps)
    uint64_t* mma_bar = full_bar + 2;              // phase-0 MMAs have finished reading the tiles
    uint64_t* tmem_full_bar = full_bar + 3;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(full_bar + 4);

    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    const int ky = blockIdx.x, parity = blockIdx.y;
    const int n0 = int(blockIdx.z) * kIpc;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        if (kMode) prefetch_tmap(&tmap_x2);
        if (!a.dz_chw) prefetch_tmap(&tmap_dz);
        mbar_init(full_bar, 1);
        mbar_init(b_bar, 4);
        mbar_init(mma_bar, 1);
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<32>(tmem_slot);
    // K rows 196..199 of every tile are never written by TMA or the repack: zero them once.
    for (int i = threadIdx.x; i < 3 * 128; i += kWgThreads)
        *reinterpret_cast<float*>(smem + (i / 128) * kTileBytes + kBoxBytes + (i % 128) * 4) = 0.0f;
    fence_proxy_async_smem();
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (threadIdx.x == 0) WG_STAMP(0);
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x == 0) WG_STAMP(1);

    if (warp == 0) {
        if (lane == 0) {
            for (int h = 0; h < 2; h++) {
                if (h == 1) { mbar_wait(mma_bar, 0); WG_STAMP(4); }
                mbar_arrive_expect_tx(full_bar, uint32_t((a.dz_chw ? 2 : 3) * kBoxBytes));
                const int img = n0 + kPhaseImgs * h;
                if (!kMode) {
                    tma_load_4d(a_tile, &tmap_x, full_bar, 0, -1, ky - 1, img);
                    tma_load_4d(a_tile + kTileBytes, &tmap_x, full_bar, 0, 0, ky - 1, img);
                    if (!a.dz_chw) tma_load_4d(b_tile, &tmap_dz, full_bar, 0, 0, 0, img);
                } else {
                    if (parity == 0) {
                        tma_load_4d(a_tile, &tmap_x2, full_bar, 0, -1, ky - 1, img);
                        tma_load_4d(a_tile + kTileBytes, &tmap_x, full_bar, 0, 0, ky - 1, img);
                    } else {
                        tma_load_4d(a_tile, &tmap_x, full_bar, 0, 0, ky - 1, img);
                        tma_load_4d(a_tile + kTileBytes, &tmap_x2, full_bar, 0, 0, ky - 1, img);
                    }
                    tma_load_5d(b_tile, &tmap_dz, full_bar, 0, parity, 0, 0, img);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(64, kOC, true, true);
            const uint32_t a_addr = smem_u32(a_tile), b_addr = smem_u32(b_tile);
            for (int h = 0; h < 2; h++) {
                mbar_wait(full_bar, uint32_t(h));
                WG_STAMP(h == 0 ? 2 : 5);
                if (a.dz_chw) mbar_wait(b_bar, uint32_t(h));
                WG_STAMP(h == 0 ? 3 : 6);
                tcgen05_fence_after();
#pragma unroll 5
                for (int kk = 0; kk < kSteps; kk++) {
                    // M-major A: the two 32-float groups are kTileBytes apart; 8 K rows further = 1024 bytes.
                    const uint64_t da = make_smem_desc(a_addr + kk * 1024, kTileBytes, 512, kSw128Base32);
                    const uint64_t db = make_smem_desc(b_addr + kk * 1024, 4096, 512, kSw128Base32);
                    umma_tf32(tmem_base, da, db, idesc, (h | kk) != 0);
                }
                umma_commit(h == 0 ? mma_bar : tmem_full_bar);
            }
        }
    } else {
        const int w = warp - 2, oc = lane, t = threadIdx.x - 64;
        float s = 0.0f;
        if (kMode == 0 && a.dz_chw) {  // (stride 1 never gets CHW gradients: the branch and its registers compile away)
            // One image at a time: coalesced float4 loads (the next image's already in flight) -> raw[oc][49] ->
            // B[pixel][oc] rows of 128 bytes in the 128B-swizzle-with-32B-atoms layout (32-byte chunk ^= row & 3).
            // Lane = oc on both sides of the transpose: raw reads are 49 floats apart (odd: no bank conflict), tile
            // writes hit 32 different banks.
            constexpr int kQuads = kImgFloats / 4;  // 392
            float4 cur[4], nxt[4];
            const float4* src0 = reinterpret_cast<const float4*>(a.dz_chw + size_t(n0) * kImgFloats);
#pragma unroll
            for (int u = 0; u < 4; u++) cur[u] = (t + 128 * u < kQuads) ? __ldg(src0 + t + 128 * u) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
            for (int i = 0; i < kIpc; i++) {
                const int g = i % kPhaseImgs;
                if (i == kPhaseImgs) mbar_wait(mma_bar, 0);  // phase-0 MMAs are done with the tile
                if (i + 1 < kIpc) {
                    const float4* src = src0 + size_t(i + 1) * kQuads;
#pragma unroll
                    for (int u = 0; u < 4; u++) nxt[u] = (t + 128 * u < kQuads) ? __ldg(src + t + 128 * u) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < 4; u++)
                    if (t + 128 * u < kQuads) reinterpret_cast<float4*>(raw)[t + 128 * u] = cur[u];
                epilogue_bar_sync();
#pragma unroll
                for (int pp = 0; pp < 13; pp++) {
                    const int p = w + 4 * pp;
                    if (p < kPix) {
                        const float v = raw[oc * kPix + p];
                        const int r = g * kPix + p;
                        *reinterpret_cast<float*>(b_tile + r * 128 + (((oc >> 3) ^ (r & 3)) << 5) + (oc & 7) * 4) = v;
                        s += v;
                    }
                }
                epilogue_bar_sync();  // raw may be overwritten
                if (g == kPhaseImgs - 1) {
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(b_bar);
                }
#pragma unroll
                for (int u = 0; u < 4; u++) cur[u] = nxt[u];
            }
        } else if (ky == 0 && parity == 0) {
            // NHWC: bias gradient straight from global memory (coalesced rows of 32 oc)
            const float* src = a.dz_nhwc + size_t(n0) * kRowsPerImg * kOC + oc;
            const int rmax = (a.n - n0) * kRowsPerImg;  // (stride 1: the last CTA may own fewer than kIpc images)
#pragma unroll 14
            for (int r = w * (kIpc * kRowsPerImg / 4); r < (w + 1) * (kIpc * kRowsPerImg / 4); r++)
                if (r < rmax) s += __ldg(src + size_t(r) * kOC);
        }
        dbp[w * 32 + oc] = s;
    }
    // Accumulator row of this thread: with M=64 each TMEM lane quarter holds 16 rows in its first 16 lanes.
    if (warp >= 2) {
        const int q = warp & 3;
        const int m = 16 * q + lane;
        epilogue_bar_sync();  // dbp complete
        mbar_wait(tmem_full_bar, 0);
        if (threadIdx.x == 64) WG_STAMP(7);
        tcgen05_fence_after();
        float v[32], t16[16];
        tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16), t16);
#pragma unroll
        for (int i = 0; i < 16; i++) v[i] = t16[i];
        tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16) + 16u, t16);
#pragma unroll
        for (int i = 0; i < 16; i++) v[16 + i] = t16[i];
        if (lane < 16) {
            if (m == kBiasRow && (kMode == 0 || parity == 0)) {  // a row without a weight in it carries this CTA's bias-gradient partial
#pragma unroll
                for (int i = 0; i < 32; i++) v[i] = (dbp[i] + dbp[32 + i]) + (dbp[64 + i] + dbp[96 + i]);
            }
            float4* out = reinterpret_cast<float4*>(a.part + (((size_t(ky) * gridDim.y + parity) * gridDim.z + blockIdx.z) * 64 + m) * 32);
#pragma unroll
            for (int i = 0; i < 8; i++) out[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
        if (threadIdx.x == 64) WG_STAMP(8);
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<32>(tmem_base);
    }
}

// dW / db = sum over the K splits of the partial tiles, in split order (fixed: results are run-to-run identical).
// One thread per 4 output channels of one (ky, kx, c) row; all of a thread's loads are independent.
__global__ void __launch_bounds__(128)
wgrad_reduce_kernel(const float* __restrict__ part, int splits, float* __restrict__ dw, float* __restrict__ db, int accumulate) {
    pdl_wait();
    pdl_launch_dependents();
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    constexpr int kItems = 3 * 48 * 8;  // + 8 bias items of (ky 0, row 0)
    if (t >= kItems + 8) return;
    const bool is_bias = t >= kItems;
    const int tt = is_bias ? t - kItems : t;
    const int ky = is_bias ? 0 : tt / 384, row = is_bias ? 0 : 16 + (tt % 384) / 8, oc4 = tt % 8;
    const float4* src = reinterpret_cast<const float4*>(part + (size_t(ky) * splits * 64 + row) * 32) + oc4;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
    for (int z = 0; z < splits; z++) {
        const float4 p4 = __ldcg(src + size_t(z) * (64 * 32 / 4));
        s.x += p4.x; s.y += p4.y; s.z += p4.z; s.w += p4.w;
    }
    const float sv[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int oc = oc4 * 4 + i;
        float* dst = is_bias ? db + oc : dw + (size_t(oc) * kC + (row & 15)) * 9 + ky * 3 + (row / 16 - 1);
        *dst = accumulate ? *dst + sv[i] : sv[i];
    }
}

// Stride 1: dW / db from the [3 ky][2 parities][splits] partial tiles.  One WARP per item (4 output channels of one
// (ky, kx, c), or of the bias): the lanes split the (parity, split) tiles, every load is in flight at once, and a shuffle
// tree in a fixed order finishes the sum (run-to-run identical).
__global__ void __launch_bounds__(128)
wgrad_s1_reduce_kernel(const float* __restrict__ part, int splits, float* __restrict__ dw, float* __restrict__ db, int accumulate) {
    pdl_wait();
    pdl_launch_dependents();
    const int item = (blockIdx.x * blockDim.x + threadIdx.x) / 32, lane = threadIdx.x % 32;
    constexpr int kItems = 3 * 48 * 8;  // + 8 bias items
    if (item >= kItems + 8) return;
    const bool is_bias = item >= kItems;
    const int tt = is_bias ? item - kItems : item;
    const int ky = is_bias ? 0 : tt / 384, kx = (tt % 384) / 128, c = (tt % 128) / 8, oc4 = tt % 8;
    // tile rows: even outputs [kx0 | kx1'] [kx1 | kx2], odd outputs [kx0 | kx1] [kx1' | kx2]; bias: row 16 of parity 0
    const int row0 = is_bias ? 16 : (kx == 0 ? 0 : kx == 1 ? 32 : 48) + c, row1 = (kx == 0 ? 0 : kx == 1 ? 16 : 48) + c;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    const int total = (is_bias ? 1 : 2) * splits;
    for (int j0 = lane; j0 < total; j0 += 4 * 32) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int j = j0 + 32 * u;
            const int par = j / splits, z = j % splits;
            v[u] = j < total ? __ldcg(reinterpret_cast<const float4*>(part + (((size_t(ky) * 2 + par) * splits + z) * 64 + (par ? row1 : row0)) * 32) + oc4)
                             : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s.x += __shfl_xor_sync(0xffffffffu, s.x, o);
        s.y += __shfl_xor_sync(0xffffffffu, s.y, o);
        s.z += __shfl_xor_sync(0xffffffffu, s.z, o);
        s.w += __shfl_xor_sync(0xffffffffu, s.w, o);
    }
    if (lane != 0) return;
    const float sv[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int oc = oc4 * 4 + i;
        float* dst = is_bias ? db + oc : dw + (size_t(oc) * kC + c) * 9 + ky * 3 + kx;
        *dst = accumulate ? *dst + sv[i] : sv[i];
    }
}

}  // namespace

static long long* g_wg_dbg = nullptr;
// Instrumentation read-back (tools/diag_wgrad.py): 16 clock64 stamps per instrumented CTA of the last launch.
extern "C" int b200cnn_debug_wgrad_stamps(long long* out, int count) {
    if (!g_wg_dbg || count > 64 * 16) return 1;
    return cudaMemcpy(out, g_wg_dbg, size_t(count) * sizeof(long long), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : 2;
}

bool tc_conv3x3s2_bwd_weight_supported(const b200cnn_conv_shape* s) {
    return s->k == 3 && s->stride == 2 && s->pad == 1 && s->c == kC && s->oc == kOC && s->oh == 7 && s->ow == 7 && s->h == 14 &&
           s->w == 14 && s->n >= kIpc && s->n % kIpc == 0;
}

size_t tc_conv3x3s2_bwd_weight_scratch_floats(int n) {
    return size_t(3) * size_t(ceil_div(n, kIpc)) * kPartFloats;
}

// x: NHWC [n][14][14][16]; dz: NHWC [n][7][7][32] or CHW [n][32][7][7]; part: scratch_floats(n) floats;
int tc_conv3x3s2_bwd_weight(const float* x_nhwc, const float* dz, int dz_is_chw, float* dw, float* db, int accumulate,
                            float* part, const b200cnn_conv_shape* s, cudaStream_t st) {
    if (!tc_conv3x3s2_bwd_weight_supported(s)) return set_error(kInvalidValue, "tc_conv3x3s2_bwd_weight: unsupported shape");
    CUtensorMap tmx, tmdz;
    {
        // pairs of horizontally adjacent pixels: (32 floats, W/2, H, N)
        const uint64_t dims[4] = {32, uint64_t(s->w / 2), uint64_t(s->h), uint64_t(s->n)};
        const uint64_t strides[4] = {0, 128, uint64_t(s->w) * kC * 4, uint64_t(s->h) * s->w * kC * 4};
        const uint32_t box[4] = {32, 7, 13, uint32_t(kPhaseImgs)};
        const uint32_t estr[4] = {1, 1, 2, 1};
        if (int rc = encode_tmap_f32(&tmx, x_nhwc, 4, dims, strides, box, estr, kTmaSw128Atom32)) return rc;
    }
    {
        // (any valid map when dz is CHW: the kernel does not touch it)
        const float* base = dz_is_chw ? x_nhwc : dz;
        const uint64_t dims[4] = {32, 7, 7, uint64_t(s->n)};
        const uint64_t strides[4] = {0, 128, 7 * 128, 49 * 128};
        const uint32_t box[4] = {32, 7, 7, uint32_t(kPhaseImgs)};
        if (int rc = encode_tmap_f32(&tmdz, base, 4, dims, strides, box, nullptr, kTmaSw128Atom32)) return rc;
    }
    const int splits = s->n / kIpc;
    static long long* dbg = [] {
        long long* p = nullptr;
        const char* v = getenv("B200CNN_WG_DEBUG");
        if (v && v[0] == '1' && cudaMalloc(&p, 64 * 16 * sizeof(long long)) == cudaSuccess) cudaMemset(p, 0, 64 * 16 * sizeof(long long));
        return p;
    }();
    g_wg_dbg = dbg;
    WgradArgs a{dz_is_chw ? dz : nullptr, dz_is_chw ? nullptr : dz, part, dw, db, s->n, dbg};
    auto kern = tc_conv3x3_bwd_weight_kernel<0>;
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kWgSmem)));
        configured = true;
    }
    B200CNN_CUDA(launch_kernel(kern, dim3(3, 1, splits), dim3(kWgThreads), kWgSmem, st, tmx, tmx, tmdz, a));
    B200CNN_CUDA(launch_kernel(wgrad_reduce_kernel, dim3(ceil_div(3 * 48 * 8 + 8, 128)), dim3(128), 0, st, static_cast<const float*>(part), splits, dw, db, accumulate));
    B200CNN_CHECK_LAUNCH("tc_conv3x3s2_bwd_weight");
    return 0;
}

bool tc_conv3x3s1_bwd_weight_supported(const b200cnn_conv_shape* s) {
    return s->k == 3 && s->stride == 1 && s->pad == 1 && s->c == kC && s->oc == kOC && s->oh == 14 && s->ow == 14 && s->h == 14 &&
           s->w == 14 && s->n >= 1;
}

size_t tc_conv3x3s1_bwd_weight_scratch_floats(int n) { return size_t(3) * 2 * size_t(ceil_div(n, 4)) * kPartFloats; }

// x: NHWC [n][14][14][16]; dz: NHWC [n][14][14][32]; part: scratch_floats(n) floats.
int tc_conv3x3s1_bwd_weight(const float* x_nhwc, const float* dz_nhwc, float* dw, float* db, int accumulate, float* part,
                            const b200cnn_conv_shape* s, cudaStream_t st) {
    if (!tc_conv3x3s1_bwd_weight_supported(s)) return set_error(kInvalidValue, "tc_conv3x3s1_bwd_weight: unsupported shape");
    CUtensorMap tme, tmo, tmdz;
    const uint64_t row = uint64_t(s->w) * kC * 4, img = uint64_t(s->h) * row;
    {
        // even-aligned pairs of horizontally adjacent pixels: (32 floats, 7 pairs, H, N) ...
        const uint64_t dims[4] = {32, 7, uint64_t(s->h), uint64_t(s->n)};
        const uint64_t strides[4] = {0, 128, row, img};
        const uint32_t box[4] = {32, 7, 14, 2};
        if (int rc = encode_tmap_f32(&tme, x_nhwc, 4, dims, strides, box, nullptr, kTmaSw128Atom32)) return rc;
        // ... and the odd-aligned ones: one pixel further, 6 pairs per row (pair 6 would reach into the next row: it is the padding)
        const uint64_t dims_o[4] = {32, 6, uint64_t(s->h), uint64_t(s->n)};
        if (int rc = encode_tmap_f32(&tmo, x_nhwc + kC, 4, dims_o, strides, box, nullptr, kTmaSw128Atom32)) return rc;
    }
    {
        const uint64_t dims[5] = {32, 2, 7, 14, uint64_t(s->n)};
        const uint64_t strides[5] = {0, 128, 256, 14 * 128, 196 * 128};
        const uint32_t box[5] = {32, 1, 7, 14, 2};
        if (int rc = encode_tmap_f32(&tmdz, dz_nhwc, 5, dims, strides, box, nullptr, kTmaSw128Atom32)) return rc;
    }
    const int splits = ceil_div(s->n, 4);
    WgradArgs a{nullptr, dz_nhwc, part, dw, db, s->n, nullptr};
    auto kern = tc_conv3x3_bwd_weight_kernel<1>;
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kWgSmem)));
        configured = true;
    }
    B200CNN_CUDA(launch_kernel(kern, dim3(3, 2, splits), dim3(kWgThreads), kWgSmem, st, tme, tmo, tmdz, a));
    B200CNN_CUDA(launch_kernel(wgrad_s1_reduce_kernel, dim3(ceil_div((3 * 48 * 8 + 8) * 32, 128)), dim3(128), 0, st, static_cast<const float*>(part),
                               splits, dw, db, accumulate));
    B200CNN_CHECK_LAUNCH("tc_conv3x3s1_bwd_weight");
    return 0;
}

}  // namespace b200cnn
