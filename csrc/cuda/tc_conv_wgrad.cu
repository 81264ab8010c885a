// Weight gradient of the 3x3 / stride-2 / pad-1 convolution (16 -> 32 channels, 14x14 -> 7x7) on the tensor
// cores, as a split-K implicit GEMM:
//     dW[oc][c][ky][kx] = sum over (n, oy, ox) of dz[n][oy][ox][oc] * x[n][2oy-1+ky][2ox-1+kx][c]
// i.e. D[M = (kx, c)][N = oc] = sum_K A[K][M] * B[K][N] with K = output pixels, for each ky.  Both operands are
// "MN-major" (the K index is the ROW of the staged tile), which is exactly what TMA delivers without any im2col:
//   A  x is NHWC with 16 channels, so two horizontally adjacent pixels are 32 contiguous floats = one 128-byte
//      row.  Seen as dims (32 floats, W/2 pairs, H, N) a box (32, 7, 13 walked with stride 2, 8 images) started
//      at pair 0 / row ky-1 is [392 pixels][taps kx=1 | kx=2], and started at pair -1 it is [392 pixels]
//      [unused | tap kx=0]; out-of-bounds pairs / rows are the zero padding.  Two boxes = one 64-row M extent.
//   B  dz [392 pixels][32 oc]: one TMA box when dz is NHWC; when it arrives CHW (what the FC layer above
//      writes) the epilogue warps transpose it into the same swizzled tile while the A boxes are in flight.
// 8 images per CTA = 392 K rows = 49 tcgen05.mma (M=128 of which 64 rows are used, N=32, K=8) into one TMEM
// tile.  Grid (3 ky, 1, n/8) in clusters of 8 along z: the 7 peers ship their 64x32 partial tiles into the
// leader's shared memory (DSMEM), the leader adds them in rank order and writes one partial per cluster; the
// last leader to finish (device counter) adds the clusters' partials in index order and scatters dW / db in
// the reference's OIHW layout.  Fixed summation order everywhere: results are run-to-run identical.
// Bias gradient: column sums of B, carried in row 0 of the partial tile (an unused tap row).
//
// Replaces the weight half of Layer_feedBack_conv (/root/reference/cnn.c:236,242) for the second convolution.
#include <cstdlib>

#include "backend.cuh"
#include "common.cuh"
#include "tc_common.cuh"

namespace b200cnn {

namespace {

using namespace tc;

constexpr int kWgThreads = 192;               // warp 0: TMA, warp 1: TMEM + MMA, warps 2-5: dz staging + epilogue
constexpr int kIpc = 8;                       // images per CTA
constexpr int kPix = 49;                      // 7 x 7 output pixels
constexpr int kRows = kIpc * kPix;            // 392 K rows = 49 MMA steps of 8
constexpr int kGroupBytes = kRows * 128;      // one [392][32 floats] tile
constexpr int kCluster = 8;
constexpr int kPartFloats = 64 * 32;          // partial tile: rows (kx+1)*16 + c, row 0 = bias gradient
constexpr int kOC = 32, kC = 16;
// The M=128 instruction shape reads four 32-float M groups kGroupBytes apart: the allocation must cover all
// four even though only two hold data (groups 2 and 3 produce accumulator rows nobody reads).  The M=64 shape reads
// two, so it fits next to other kernels' CTAs on the same SM.
template <bool M64>
constexpr size_t wg_smem() { return (M64 ? 3 : 4) * size_t(kGroupBytes) + 4 * 32 * 4 + 64 + 1024; }
static_assert((kCluster - 1) * kPartFloats * 4 <= 2 * kGroupBytes, "peers' partial tiles reuse the A tiles");

struct WgradArgs {
    const float* dz_chw;  // non-NULL: dz as [n][32][49]; NULL: NHWC through tmap_dz
    float* part;          // [3][clusters][64][32]
    unsigned* counter;    // zero before the first launch; resets itself
    float* dw;            // OIHW [32][16][3][3]
    float* db;            // [32]
    int n, clusters, accumulate;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void epilogue_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

template <bool M64>
__global__ void __launch_bounds__(kWgThreads, 1)
tc_conv3x3s2_bwd_weight_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_dz,
                               WgradArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_tile = smem;                        // 2 groups: [unused | kx0], [kx1 | kx2]
    uint8_t* b_tile = smem + 2 * kGroupBytes;      // dz
    float* red = reinterpret_cast<float*>(smem);   // leader, after the MMAs: [7][64][32] partial tiles of the peers
    float* dbp = reinterpret_cast<float*>(smem + (M64 ? 3 : 4) * kGroupBytes);  // [4 warps][32] bias-gradient partials
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(dbp + 4 * 32);
    uint64_t* tmem_full_bar = full_bar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
    uint32_t* last_flag = tmem_slot + 1;

    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    const int ky = blockIdx.x;
    const int rank = int(cluster_ctarank());
    const int cl = int(blockIdx.z) / kCluster;
    const int n0 = int(blockIdx.z) * kIpc;
    const bool has = n0 < a.n;  // (n is a multiple of 8; trailing CTAs of the last cluster may have nothing)

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        if (!a.dz_chw) prefetch_tmap(&tmap_dz);
        mbar_init(full_bar, 1);
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<32>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0 && has) {
            mbar_arrive_expect_tx(full_bar, uint32_t((a.dz_chw ? 2 : 3) * kGroupBytes));
            tma_load_4d(a_tile, &tmap_x, full_bar, 0, -1, ky - 1, n0);
            tma_load_4d(a_tile + kGroupBytes, &tmap_x, full_bar, 0, 0, ky - 1, n0);
            if (!a.dz_chw) tma_load_4d(b_tile, &tmap_dz, full_bar, 0, 0, 0, n0);
        }
    } else if (warp >= 2 && a.dz_chw) {
        // CHW -> [pixel][oc] rows of 128 bytes in the 128B-swizzle-with-32B-atoms layout (32-byte chunk index
        // ^= row & 3).  Lane = oc: the 32 lanes of a store hit 32 different banks; each lane reads its 49
        // contiguous floats of (image, oc) with all loads in flight.
        const int w = warp - 2, oc = lane;
        float s = 0.0f;
        if (has) {
#pragma unroll 1
            for (int gi = 0; gi < kIpc / 4; gi++) {
                const int g = w + 4 * gi;
                const float* src = a.dz_chw + (size_t(n0 + g) * kOC + oc) * kPix;
                float v[kPix];
#pragma unroll
                for (int p = 0; p < kPix; p++) v[p] = __ldg(src + p);
#pragma unroll
                for (int p = 0; p < kPix; p++) {
                    const int r = g * kPix + p;
                    *reinterpret_cast<float*>(b_tile + r * 128 + (((oc >> 3) ^ (r & 3)) << 5) + (oc & 7) * 4) = v[p];
                    s += v[p];
                }
            }
            fence_proxy_async_smem();
        }
        dbp[w * 32 + oc] = s;
    }
    __syncthreads();

    if (warp == 1) {
        if (lane == 0 && has) {
            constexpr uint32_t idesc = make_idesc_tf32(M64 ? 64 : 128, kOC, true, true);
            mbar_wait(full_bar, 0);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_tile), b_addr = smem_u32(b_tile);
#pragma unroll 7
            for (int kk = 0; kk < kRows / 8; kk++) {
                // M-major A: 32-float groups kGroupBytes apart (groups 2, 3 of the M=128 shape read whatever follows:
                // their accumulator rows 64..127 are never looked at).  8 K rows further = 1024 bytes.
                const uint64_t da = make_smem_desc(a_addr + kk * 1024, kGroupBytes, 512, kSw128Base32);
                const uint64_t db = make_smem_desc(b_addr + kk * 1024, 4096, 512, kSw128Base32);
                umma_tf32(tmem_base, da, db, idesc, kk != 0);
            }
            umma_commit(tmem_full_bar);
        }
    }
    float v[32];
    const int q = warp & 3;
    // Accumulator row of this thread.  M=128: row = TMEM lane (rows 0..63 live in lane quarters 0 and 1).
    // M=64: each lane quarter holds 16 rows in its first 16 lanes.
    const int m = M64 ? 16 * q + lane : 32 * q + lane;
    const bool ld_warp = warp >= 2 && (M64 || q < 2);
    const bool epi = ld_warp && (!M64 || lane < 16);
    if (warp >= 2) {
        if (!a.dz_chw) {
            // NHWC: bias gradient = column sums of the staged dz tile (only the ky = 0 CTAs report it)
            const int w = warp - 2, oc = lane;
            float s = 0.0f;
            if (has && ky == 0) {
                mbar_wait(full_bar, 0);
#pragma unroll 14
                for (int r = w * (kRows / 4); r < (w + 1) * (kRows / 4); r++)
                    s += *reinterpret_cast<const float*>(b_tile + r * 128 + (((oc >> 3) ^ (r & 3)) << 5) + (oc & 7) * 4);
            }
            dbp[w * 32 + oc] = s;
        }
        epilogue_bar_sync();  // dbp complete
        if (ld_warp) {
            if (has) {
                mbar_wait(tmem_full_bar, 0);
                tcgen05_fence_after();
                float t[16];
                tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16), t);
#pragma unroll
                for (int i = 0; i < 16; i++) v[i] = t[i];
                tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16) + 16u, t);
#pragma unroll
                for (int i = 0; i < 16; i++) v[16 + i] = t[i];
            } else {
#pragma unroll
                for (int i = 0; i < 32; i++) v[i] = 0.0f;
            }
            if (epi && m == 0) {
#pragma unroll
                for (int i = 0; i < 32; i++) v[i] = (dbp[i] + dbp[32 + i]) + (dbp[64 + i] + dbp[96 + i]);
            }
        }
    }
    // Everyone's MMAs have finished reading shared memory: the leader's A tiles may now receive the partial tiles.
    tcgen05_fence_before();
    cluster_sync_all();
    if (epi && rank != 0) {
        const uint32_t dst = map_to_cluster_rank(smem_u32(red + (size_t(rank - 1) * 64 + m) * 32), 0);
#pragma unroll
        for (int i = 0; i < 8; i++) st_cluster_f32x4(dst + 16 * i, v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
    cluster_sync_all();
    if (rank == 0) {
        if (epi) {
#pragma unroll 1
            for (int p = 0; p < kCluster - 1; p++) {
                const float4* pr = reinterpret_cast<const float4*>(red + (size_t(p) * 64 + m) * 32);
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const float4 t4 = pr[i];
                    v[4 * i] += t4.x; v[4 * i + 1] += t4.y; v[4 * i + 2] += t4.z; v[4 * i + 3] += t4.w;
                }
            }
            float4* out = reinterpret_cast<float4*>(a.part + ((size_t(ky) * a.clusters + cl) * 64 + m) * 32);
#pragma unroll
            for (int i = 0; i < 8; i++) out[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            __threadfence();
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned total = 3u * unsigned(a.clusters);
            *last_flag = (atomicAdd(a.counter, 1u) == total - 1) ? 1u : 0u;
        }
        __syncthreads();
        if (*last_flag) {
            __threadfence();
            // (ky, row 16..63, 4 oc) items + the 8 bias items of (ky 0, row 0)
            const int items = 3 * 48 * 8 + 8;
            for (int t = threadIdx.x; t < items; t += kWgThreads) {
                const bool is_bias = t >= 3 * 48 * 8;
                const int tt = is_bias ? t - 3 * 48 * 8 : t;
                const int kyy = is_bias ? 0 : tt / 384, row = is_bias ? 0 : 16 + (tt % 384) / 8, oc4 = tt % 8;
                float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int c2 = 0; c2 < a.clusters; c2++) {
                    const float4 p4 = __ldcg(reinterpret_cast<const float4*>(a.part + ((size_t(kyy) * a.clusters + c2) * 64 + row) * 32) + oc4);
                    s.x += p4.x; s.y += p4.y; s.z += p4.z; s.w += p4.w;
                }
                const float sv[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const int oc = oc4 * 4 + i;
                    float* dst = is_bias ? a.db + oc : a.dw + (size_t(oc) * kC + (row & 15)) * 9 + kyy * 3 + (row / 16 - 1);
                    *dst = a.accumulate ? *dst + sv[i] : sv[i];
                }
            }
            if (threadIdx.x == 0) *a.counter = 0;
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<32>(tmem_base);
    }
}

}  // namespace

bool tc_conv3x3s2_bwd_weight_supported(const b200cnn_conv_shape* s) {
    return s->k == 3 && s->stride == 2 && s->pad == 1 && s->c == kC && s->oc == kOC && s->oh == 7 && s->ow == 7 && s->h == 14 &&
           s->w == 14 && s->n >= kIpc && s->n % kIpc == 0;
}

size_t tc_conv3x3s2_bwd_weight_scratch_floats(int n) {
    return size_t(3) * size_t(ceil_div(n, kIpc * kCluster)) * kPartFloats;
}

// x: NHWC [n][14][14][16]; dz: NHWC [n][7][7][32] or CHW [n][32][7][7]; part: scratch_floats(n) floats;
// counter: one zero-initialised word (the kernel leaves it zero).
int tc_conv3x3s2_bwd_weight(const float* x_nhwc, const float* dz, int dz_is_chw, float* dw, float* db, int accumulate,
                            float* part, unsigned* counter, const b200cnn_conv_shape* s, cudaStream_t st) {
    if (!tc_conv3x3s2_bwd_weight_supported(s)) return set_error(kInvalidValue, "tc_conv3x3s2_bwd_weight: unsupported shape");
    CUtensorMap tmx, tmdz;
    {
        // pairs of horizontally adjacent pixels: (32 floats, W/2, H, N)
        const uint64_t dims[4] = {32, uint64_t(s->w / 2), uint64_t(s->h), uint64_t(s->n)};
        const uint64_t strides[4] = {0, 128, uint64_t(s->w) * kC * 4, uint64_t(s->h) * s->w * kC * 4};
        const uint32_t box[4] = {32, 7, 13, uint32_t(kIpc)};
        const uint32_t estr[4] = {1, 1, 2, 1};
        if (int rc = encode_tmap_f32(&tmx, x_nhwc, 4, dims, strides, box, estr, kTmaSw128Atom32)) return rc;
    }
    {
        // (any valid map when dz is CHW: the kernel does not touch it)
        const float* base = dz_is_chw ? x_nhwc : dz;
        const uint64_t dims[4] = {32, 7, 7, uint64_t(s->n)};
        const uint64_t strides[4] = {0, 128, 7 * 128, 49 * 128};
        const uint32_t box[4] = {32, 7, 7, uint32_t(kIpc)};
        if (int rc = encode_tmap_f32(&tmdz, base, 4, dims, strides, box, nullptr, kTmaSw128Atom32)) return rc;
    }
    const int clusters = ceil_div(s->n, kIpc * kCluster);
    WgradArgs a{dz_is_chw ? dz : nullptr, part, counter, dw, db, s->n, clusters, accumulate};
    static const bool m64 = [] {
        const char* v = getenv("B200CNN_WGRAD_M64");
        return !(v && v[0] == '0');  // default: the M=64 shape (smaller footprint); 0 = M=128
    }();
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(tc_conv3x3s2_bwd_weight_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(wg_smem<true>())));
        B200CNN_CUDA(cudaFuncSetAttribute(tc_conv3x3s2_bwd_weight_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(wg_smem<false>())));
        configured = true;
    }
    const dim3 grid(3, 1, clusters * kCluster);
    if (m64)
        B200CNN_CUDA(launch_kernel_cluster(tc_conv3x3s2_bwd_weight_kernel<true>, grid, dim3(kWgThreads), wg_smem<true>(), st, kCluster, tmx, tmdz, a));
    else
        B200CNN_CUDA(launch_kernel_cluster(tc_conv3x3s2_bwd_weight_kernel<false>, grid, dim3(kWgThreads), wg_smem<false>(), st, kCluster, tmx, tmdz, a));
    B200CNN_CHECK_LAUNCH("tc_conv3x3s2_bwd_weight");
    return 0;
}

}  // namespace b200cnn
