// The fully connected part of a training step — forward AND the error-propagation half of backward — as ONE
// cluster kernel instead of seven dependent launches:
//     h1  = tanh(x W1^T + b1)          h2 = tanh(h1 W2^T + b2)          p = softmax(h2 W3^T + b3)
//     dz3 = p - onehot (+ the reference's metrics)                       dz2 = (dz3 W3) * (1 - h2^2)
//     dz1 = (dz2 W2) * (1 - h1^2)      dx = (dz1 W1) * act'(x)
// At batch 256 every one of these GEMMs is latency-bound, so a step was a chain of kernel boundaries; the samples of
// a mini-batch, however, never interact in these seven operations.  So a CLUSTER of 8 CTAs owns 16 samples and walks
// the whole chain for them, exchanging activations through distributed shared memory:
//   * the GEMMs are computed transposed, D[feature][sample] = W * act^T: features are the M dimension of
//     tcgen05.mma (rows -> TMEM lanes), the 16 samples are N.  All operands are the row-major [out][in] tensors the
//     reference stores (/root/reference/cnn.c:116-121): K-major A for the forward GEMMs, M-major A (the transposed
//     view of the same memory) for the backward ones — no transposed copies exist;
//   * measured on B200: a tcgen05.mma this small costs 60-100 ns whatever its shape (it does not overlap its
//     neighbours), so the number of MMAs a CTA issues back to back IS its GEMM time.  GEMM 1 (K = 1568) is therefore
//     split over K: every CTA multiplies all n1 features by its eighth of K (<= 56 MMAs instead of 196), the partial
//     [n1 x 16] tiles are reduce-scattered through DSMEM (13 KB per CTA) and summed in CTA order (deterministic).
//     The last GEMM (M = 1568 outputs) is split over M: two M = 128 tiles per CTA, 50 MMAs.  The four GEMMs in the
//     middle give every CTA a 28-feature slice (M = 64 tile, 25 MMAs);
//   * at a layer boundary every CTA stores its slice of the new activations straight into ALL eight CTAs' next B
//     operand tile (st.shared::cluster, already in the swizzled K-major layout), fences the async proxy, and
//     arrives on every CTA's mbarrier with release at cluster scope; the MMA thread of each CTA waits on its own
//     barrier.  The same values go to global memory for the weight-gradient kernels;
//   * W2 / W3 and their transposed views are prefetched into shared memory while GEMM 1 runs; W1 streams through a
//     4-stage TMA ring twice (K-major for GEMM 1, M-major for GEMM 6; the second pass prefetches under the middle);
//   * the softmax layer (<= 16 classes) is computed redundantly by every CTA, so its gradient tile is local.
// Replaces, per mini-batch, Layer_feedForw_full x3 incl. softmax (/root/reference/cnn.c:113-152), the loss gradient
// and metric (cnn.c:275-286) and the error-propagation half of Layer_feedBack_full x3 (cnn.c:158-166).
#include <cstdlib>

#include "backend.cuh"
#include "common.cuh"
#include "tc_common.cuh"

namespace b200cnn {

namespace {

using namespace tc;

constexpr int CL = 8;              // CTAs per cluster
constexpr int kThreads = 192;      // warp 0: ring producer, warp 1: TMEM + MMA issue, warps 2-5: epilogues
constexpr int kMaxKB = 7;          // K blocks over a hidden layer: hidden sizes up to 208 (two M = 128 tiles: 128 + 80)
constexpr int kMaxSlice = 28;      // features per CTA in the middle GEMMs (a multiple of 4, 8 x 28 >= 208)
constexpr int kQuads = kMaxSlice / 4;

// NB = samples per cluster = UMMA N.  A B200 holds at most 15 clusters of 8 CTAs with this much shared memory
// (cudaOccupancyMaxActiveClusters), so a batch of 256 runs as 11 clusters of 24 samples in ONE wave instead of 16
// clusters of 16 in two; smaller batches use 16.
// MID: only the four GEMMs in the middle (h1 comes from global memory, the chain ends with dz1): no W1 ring, half the
// shared memory, so two CTAs share an SM and the kernel overlaps its neighbours in the stream.
template <int NB, bool MID = false>
struct Cfg {
    static constexpr int kTileB = NB * 128;   // one K block of a B operand (NB samples x 32 floats)
    // ring stage.  GEMM 1: x tile | W1 rows 0-127 16 KiB | W1 rows 128.. <= 10 KiB;  GEMM 6: up to seven 4 KiB groups
    static constexpr int kStage = (kTileB + 16384 + 10240) > 28672 ? (kTileB + 16384 + 10240) : 28672;
    static constexpr int kStages = MID ? 0 : (NB <= 16 ? 4 : 3);
    // shared memory map (bytes from the 1 KiB-aligned base)
    static constexpr int kOffRing = 0;
    static constexpr int kOffW2k = kOffRing + kStages * kStage;   // 7 x 4 KiB  W2 rows [s2*c, +32), K-major      (GEMM 2 A)
    static constexpr int kOffW2m = kOffW2k + kMaxKB * 4096;       // 7 x 4 KiB  W2 cols [s1*c, +32), M-major      (GEMM 5 A)
    static constexpr int kOffW3k = kOffW2m + kMaxKB * 4096;       // 7 x 2 KiB  W3 (16 rows), K-major             (GEMM 3 A)
    static constexpr int kOffW3m = kOffW3k + kMaxKB * 2048;       // 2 KiB      W3 cols [s2*c, +32), M-major      (GEMM 4 A)
    static constexpr int kOffT0 = kOffW3m + 2048;                 // 7 K blocks: h1 tile, later dz2 tile          (B)
    static constexpr int kOffT1 = kOffT0 + kMaxKB * kTileB;       // 7 K blocks: GEMM-1 partials of the 8 CTAs, then h2 tile, then dz1 tile
    static constexpr int kOffZ3 = kOffT1 + kMaxKB * kTileB;       // 1 K block: dz3 tile, written by the softmax  (B)
    static constexpr int kOffH1s = kOffZ3 + kTileB;               // [28][NB] floats: this CTA's h1 slice (for 1 - h1^2)
    static constexpr int kOffH2s = kOffH1s + kMaxSlice * NB * 4;  // [28][NB] this CTA's h2 slice
    static constexpr int kOffStg = kOffH2s + kMaxSlice * NB * 4;  // [28][NB] transposition buffer of epilogues 4 and 5
    static constexpr int kOffSmx = kOffStg + kMaxSlice * NB * 4;  // [16][NB + 1] logits, class-major, for the per-sample softmax
    static constexpr int kOffBars = (kOffSmx + 16 * (NB + 1) * 4 + 63) & ~63;
    static constexpr int kNumBars = 2 * kStages + 1 + 7 + 5 + 1 + 1;
    static constexpr size_t kSmem = size_t(kOffBars) + kNumBars * 8 + 16 + 1024;
    static_assert(kSmem <= 227 * 1024, "shared memory budget");
    static_assert(CL * kMaxSlice * NB * 4 <= kMaxKB * kTileB, "GEMM-1 partials must fit the tile they borrow");
    // accumulators, 32 columns apart: GEMM 1 tiles at 0 / 32, GEMMs 2-5 at 64 / 96 / 128 / 160, GEMM 6 tiles at 192 / 224
    static constexpr int kTmemCols = 256;
};

struct ChainArgs {
    const float *b1, *b2, *b3;
    const uint8_t* labels;
    const float* x;      // [m][k1]   input activations (for act'(x) of the last GEMM)
    float *h1, *h2;      // [m][n1], [m][n2]
    float *probs, *dz3;  // [m][ldc]
    float *dz2, *dz1;    // [m][n2], [m][n1]
    float* dx;           // [m][k1]
    float* metrics;      // [4] or NULL
    int m, k1, n1, n2, nc, ldc, act_x;
    long long* stamps;   // debugging aid: SM clock of milestones, written by CTA (0, 0, 0) only
    int stop;            // debugging aid (B200CNN_CHAIN_DBG): run only the first `stop` GEMMs; 0 = none, >= 6 everything
};

#define CHAIN_STAMP(i)                                                               \
    do {                                                                             \
        if (a.stamps && blockIdx.y == 0 && blockIdx.z == 0) a.stamps[i] = clock64(); \
    } while (0)

__device__ __forceinline__ void mbar_arrive_remote(uint32_t remote_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote_addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait_cluster(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait_cluster(bar, parity)) {
        if (clock64() - t0 > 2000000000ll) {
            printf("b200cnn: cluster barrier wait timed out (block %d,%d,%d thread %d)\n", blockIdx.x, blockIdx.y, blockIdx.z, threadIdx.x);
            asm volatile("trap;");
        }
    }
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// NB consecutive accumulator columns (16 or 24) of this warp's TMEM lane quarter.
template <int NB>
__device__ __forceinline__ void tmem_ld_row(uint32_t taddr, float (&f)[NB]) {
    static_assert(NB == 16 || NB == 24, "supported cluster batch sizes");
    uint32_t r[24];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    if (NB == 24)
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23])
                     : "r"(taddr + 16u)
                     : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < NB; i++) f[i] = __uint_as_float(r[i]);
}

// Byte offset of float k of sample n inside a K-major, 128-byte-swizzled B operand made of NB x 128 byte K blocks.
template <int NB>
__device__ __forceinline__ uint32_t tile_off(int n, int k) {
    return uint32_t((k >> 5) * (NB * 128) + n * 128 + ((((k & 31) >> 2) ^ (n & 7)) << 4) + (k & 3) * 4);
}

template <int NB, bool MID>
__global__ void __launch_bounds__(kThreads, 1)
tc_fc_chain_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w1a,
                   const __grid_constant__ CUtensorMap tm_w1b, const __grid_constant__ CUtensorMap tm_w1m,
                   const __grid_constant__ CUtensorMap tm_w2k, const __grid_constant__ CUtensorMap tm_w2m,
                   const __grid_constant__ CUtensorMap tm_w3k, const __grid_constant__ CUtensorMap tm_w3m,
                   const __grid_constant__ CUtensorMap tm_h1, ChainArgs a) {
    using C = Cfg<NB, MID>;
    const int stop = MID ? (a.stop < 5 ? a.stop : 5) : a.stop;  // MID ends with dz1 (epilogue 5 stores to global memory only)
    constexpr int kStages = C::kStages, kStage = C::kStage, kTileB = C::kTileB;
    constexpr int kOffRing = C::kOffRing, kOffW2k = C::kOffW2k, kOffW2m = C::kOffW2m, kOffW3k = C::kOffW3k, kOffW3m = C::kOffW3m;
    constexpr int kOffT0 = C::kOffT0, kOffT1 = C::kOffT1, kOffZ3 = C::kOffZ3;
    constexpr int kTmemCols = C::kTmemCols;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kOffBars);
    uint64_t* full = bars;                   // [kStages]
    uint64_t* empty = full + kStages;        // [kStages]
    uint64_t* wbar = empty + kStages;        // small weights landed
    uint64_t* acc_bar = wbar + 1;            // [7] accumulators complete: GEMM 1 (both tiles), GEMMs 2-5, GEMM 6 (both tiles)
    uint64_t* xbar = acc_bar + 7;            // [5] cluster-wide: all 8 CTAs have delivered their part (0: partials, 1-4: tiles)
    uint64_t* z3_bar = xbar + 5;             // the softmax has written the dz3 tile
    uint64_t* tile_bar = z3_bar + 1;         // MID: the h1 tile has landed from global memory
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tile_bar + 1);
    float* red = reinterpret_cast<float*>(smem + kOffT1);    // [CL src][kMaxSlice][NB]
    float* h1s = reinterpret_cast<float*>(smem + C::kOffH1s);
    float* h2s = reinterpret_cast<float*>(smem + C::kOffH2s);
    float* stg = reinterpret_cast<float*>(smem + C::kOffStg);
    float* smx = reinterpret_cast<float*>(smem + C::kOffSmx);

    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    const int c = int(blockIdx.z);            // rank in the cluster (cluster = all of grid.z)
    const int n0 = int(blockIdx.y) * NB;      // first sample of this cluster
    // feature slice of a CTA in the middle GEMMs (<= 28), a multiple of 4 floats: the slice offset is the INNER
    // coordinate of the M-major TMA boxes, and a box must start on a 16-byte boundary of global memory
    const int s1 = (((a.n1 + CL - 1) / CL) + 3) & ~3, s2 = (((a.n2 + CL - 1) / CL) + 3) & ~3;
    const int sx = (((a.k1 + CL - 1) / CL) + 3) & ~3;                  // input-feature slice of the last GEMM (<= 224)
    const int kb1 = (a.k1 + 31) / 32;         // K blocks of GEMM 1, split over the cluster:
    const int kbA = kb1 * c / CL, kbB = kb1 * (c + 1) / CL;            //   this CTA's share
    const int kbh1 = (a.n1 + 31) / 32, kbh2 = (a.n2 + 31) / 32;        // K blocks over n1 / n2
    const int f1 = s1 * c, f2 = s2 * c, j0 = sx * c;                   // first feature of this CTA's slices
    const int t1 = (a.n1 + 127) / 128;        // M = 128 tiles of GEMM 1 (1 or 2)
    const int jn = min(sx, max(a.k1 - j0, 0));                         // input features of this CTA in GEMM 6
    const int xt = (jn + 127) / 128;          // M = 128 tiles of GEMM 6 (0..2)
    const int xg = (jn + 31) / 32;            // 32-feature groups to load per K block

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm_x); prefetch_tmap(&tm_w1a); prefetch_tmap(&tm_w1b); prefetch_tmap(&tm_w1m);
        prefetch_tmap(&tm_w2k); prefetch_tmap(&tm_w2m); prefetch_tmap(&tm_w3k); prefetch_tmap(&tm_w3m);
        for (int s = 0; s < kStages; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_init(wbar, 1);
        for (int i = 0; i < 7; i++) mbar_init(&acc_bar[i], 1);
        for (int i = 0; i < 5; i++) mbar_init(&xbar[i], CL);
        mbar_init(z3_bar, 1);
        mbar_init(tile_bar, 1);
        if (MID) prefetch_tmap(&tm_h1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    cluster_sync_all();  // every CTA's barriers exist before any peer arrives on them
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();               // everything above overlapped the previous kernel's tail
    pdl_launch_dependents();
    if (threadIdx.x == 0) CHAIN_STAMP(0);

    if (warp == 0) {
        // ===== ring producer: small weights, GEMM 1's share of W1 and x, then W1's M-major view for GEMM 6 =====
        if (lane == 0 && stop >= 1) {
            if constexpr (MID) {
                // small weights, then this cluster's h1 tile (all n1 features of its NB samples) as the first B operand
                mbar_arrive_expect_tx(wbar, uint32_t(kbh1 * 4096 + kbh2 * 4096 + kbh2 * 2048 + 2048));
                for (int k = 0; k < kbh1; k++) tma_load_2d(smem + kOffW2k + k * 4096, &tm_w2k, wbar, k * 32, f2);
                mbar_arrive_expect_tx(tile_bar, uint32_t(kbh1 * kTileB));
                for (int k = 0; k < kbh1; k++) tma_load_2d(smem + kOffT0 + k * kTileB, &tm_h1, tile_bar, k * 32, n0);
                for (int k = 0; k < kbh2; k++) tma_load_2d(smem + kOffW3k + k * 2048, &tm_w3k, wbar, k * 32, 0);
                tma_load_2d(smem + kOffW3m, &tm_w3m, wbar, f2, 0);
                for (int k = 0; k < kbh2; k++) tma_load_2d(smem + kOffW2m + k * 4096, &tm_w2m, wbar, f1, k * 32);
            } else {
            // GEMM 1's first stages go out first (they are the head of the critical path), then the small weights
                // (needed from GEMM 2 on), then the rest of the ring traffic as slots come free.
                int fill = 0;
                const uint32_t rows_b = uint32_t(a.n1 > 128 ? ((a.n1 - 128 + 7) & ~7) : 0);  // box rows of W1's second tile
                auto gemm1_stage = [&](int kb) {
                    const int s = fill % kStages;
                    mbar_wait(&empty[s], ((fill / kStages) & 1) ^ 1);
                    uint8_t* st = smem + kOffRing + s * kStage;
                    mbar_arrive_expect_tx(&full[s], uint32_t(kTileB + 16384 + rows_b * 128));
                    tma_load_2d(st, &tm_x, &full[s], kb * 32, n0);
                    tma_load_2d(st + kTileB, &tm_w1a, &full[s], kb * 32, 0);
                    if (rows_b) tma_load_2d(st + kTileB + 16384, &tm_w1b, &full[s], kb * 32, 128);
                    fill++;
                };
                int kb = kbA;
                for (; kb < kbB && fill < kStages; kb++) gemm1_stage(kb);
                mbar_arrive_expect_tx(wbar, uint32_t(kbh1 * 4096 + kbh2 * 4096 + kbh2 * 2048 + 2048));
                for (int k = 0; k < kbh1; k++) tma_load_2d(smem + kOffW2k + k * 4096, &tm_w2k, wbar, k * 32, f2);
                for (int k = 0; k < kbh2; k++) tma_load_2d(smem + kOffW3k + k * 2048, &tm_w3k, wbar, k * 32, 0);
                tma_load_2d(smem + kOffW3m, &tm_w3m, wbar, f2, 0);
                for (int k = 0; k < kbh2; k++) tma_load_2d(smem + kOffW2m + k * 4096, &tm_w2m, wbar, f1, k * 32);
                for (; kb < kbB; kb++) gemm1_stage(kb);
                CHAIN_STAMP(1);
                for (int kb = 0; kb < kbh1 && stop >= 6 && xt > 0; kb++, fill++) {
                    const int s = fill % kStages;
                    mbar_wait(&empty[s], ((fill / kStages) & 1) ^ 1);
                    uint8_t* st = smem + kOffRing + s * kStage;
                    mbar_arrive_expect_tx(&full[s], uint32_t(xg * 4096));
                    for (int g = 0; g < xg; g++) tma_load_2d(st + g * 4096, &tm_w1m, &full[s], j0 + 32 * g, kb * 32);
                }
            }
            CHAIN_STAMP(2);
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0 && stop >= 1) {
            constexpr uint32_t idesc_kk64 = make_idesc_tf32(64, NB, false, false);    // A K-major, B K-major, M = 64
            constexpr uint32_t idesc_mk64 = make_idesc_tf32(64, NB, true, false);     // A M-major, B K-major, M = 64
            constexpr uint32_t idesc_kk128 = make_idesc_tf32(128, NB, false, false);
            constexpr uint32_t idesc_mk128 = make_idesc_tf32(128, NB, true, false);
            const uint32_t ring = smem_u32(smem + kOffRing);
            int take = 0;
            if (!MID) {
            // GEMM 1 (this CTA's K blocks, all n1 features): partial h1^T = W1[:, K part] x^T[K part]
                for (int kb = kbA; kb < kbB; kb++, take++) {
                    const int s = take % kStages;
                    mbar_wait(&full[s], (take / kStages) & 1);
                    tcgen05_fence_after();
                    const uint32_t b_addr = ring + s * kStage, a_addr = b_addr + kTileB;
                    const int ksteps = min(4, (a.k1 - kb * 32 + 7) / 8);
                    for (int t = 0; t < t1; t++)
                        for (int kk = 0; kk < ksteps; kk++)
                            umma_tf32(tmem_base + 32 * t, make_smem_desc(a_addr + t * 16384 + kk * 32, 16, 1024, kSw128),
                                      make_smem_desc(b_addr + kk * 32, 16, 1024, kSw128), idesc_kk128, (kb != kbA) || kk != 0);
                    umma_commit(&empty[s]);
                }
                CHAIN_STAMP(3);
                umma_commit(&acc_bar[0]);
            }
            mbar_wait(wbar, 0);
            if (stop < 2) goto mma_done;
            // GEMM 2: h2^T slice = W2[f2.., :] h1^T     (boundary 1: every CTA has written its h1 slice into T0)
            if (MID) mbar_wait(tile_bar, 0);
            else mbar_wait_cluster(&xbar[1], 0);
            CHAIN_STAMP(4);
            tcgen05_fence_after();
            for (int kb = 0; kb < kbh1; kb++) {
                const int ksteps = min(4, (a.n1 - kb * 32 + 7) / 8);
                for (int kk = 0; kk < ksteps; kk++)
                    umma_tf32(tmem_base + 64, make_smem_desc(smem_u32(smem + kOffW2k + kb * 4096) + kk * 32, 16, 1024, kSw128),
                              make_smem_desc(smem_u32(smem + kOffT0 + kb * kTileB) + kk * 32, 16, 1024, kSw128), idesc_kk64, (kb | kk) != 0);
            }
            umma_commit(&acc_bar[1]);
            CHAIN_STAMP(5);
            if (stop < 3) goto mma_done;
            // GEMM 3: logits^T = W3 h2^T (all classes, every CTA)
            mbar_wait_cluster(&xbar[2], 0);
            CHAIN_STAMP(6);
            tcgen05_fence_after();
            for (int kb = 0; kb < kbh2; kb++) {
                const int ksteps = min(4, (a.n2 - kb * 32 + 7) / 8);
                for (int kk = 0; kk < ksteps; kk++)
                    umma_tf32(tmem_base + 96, make_smem_desc(smem_u32(smem + kOffW3k + kb * 2048) + kk * 32, 16, 1024, kSw128),
                              make_smem_desc(smem_u32(smem + kOffT1 + kb * kTileB) + kk * 32, 16, 1024, kSw128), idesc_kk64, (kb | kk) != 0);
            }
            umma_commit(&acc_bar[2]);
            CHAIN_STAMP(7);
            if (stop < 4) goto mma_done;
            // GEMM 4: (dz3 W3)^T slice = W3[:, f2..]^T dz3^T     (A M-major: 16 class rows x 32 features)
            mbar_wait(z3_bar, 0);
            CHAIN_STAMP(8);
            tcgen05_fence_after();
            for (int kk = 0; kk < 2; kk++)
                umma_tf32(tmem_base + 128, make_smem_desc(smem_u32(smem + kOffW3m) + kk * 1024, 4096, 512, kSw128Base32),
                          make_smem_desc(smem_u32(smem + kOffZ3) + kk * 32, 16, 1024, kSw128), idesc_mk64, kk != 0);
            umma_commit(&acc_bar[3]);
            if (stop < 5) goto mma_done;
            // GEMM 5: (dz2 W2)^T slice = W2[:, f1..]^T dz2^T
            mbar_wait_cluster(&xbar[3], 0);
            CHAIN_STAMP(9);
            tcgen05_fence_after();
            for (int kb = 0; kb < kbh2; kb++) {
                const int ksteps = min(4, (a.n2 - kb * 32 + 7) / 8);
                for (int kk = 0; kk < ksteps; kk++)
                    umma_tf32(tmem_base + 160, make_smem_desc(smem_u32(smem + kOffW2m + kb * 4096) + kk * 1024, 4096, 512, kSw128Base32),
                              make_smem_desc(smem_u32(smem + kOffT0 + kb * kTileB) + kk * 32, 16, 1024, kSw128), idesc_mk64, (kb | kk) != 0);
            }
            umma_commit(&acc_bar[4]);
            CHAIN_STAMP(10);
            if (stop < 6) goto mma_done;
            if constexpr (!MID) {
            // GEMM 6: (dz1 W1)^T slice = W1[:, j0..]^T dz1^T, up to two M = 128 tiles; W1 arrives through the ring
                mbar_wait_cluster(&xbar[4], 0);
                CHAIN_STAMP(11);
                tcgen05_fence_after();
                for (int kb = 0; kb < kbh1 && xt > 0; kb++, take++) {
                    const int s = take % kStages;
                    mbar_wait(&full[s], (take / kStages) & 1);
                    tcgen05_fence_after();
                    const int ksteps = min(4, (a.n1 - kb * 32 + 7) / 8);
                    for (int t = 0; t < xt; t++)
                        for (int kk = 0; kk < ksteps; kk++)
                            umma_tf32(tmem_base + 192 + 32 * t, make_smem_desc(ring + s * kStage + t * 16384 + kk * 1024, 4096, 512, kSw128Base32),
                                      make_smem_desc(smem_u32(smem + kOffT1 + kb * kTileB) + kk * 32, 16, 1024, kSw128), idesc_mk128, (kb | kk) != 0);
                    umma_commit(&empty[s]);
                }
                umma_commit(&acc_bar[5]);
                CHAIN_STAMP(12);
            }
        mma_done:;
        }
    } else if (stop >= 1) {
        // ===== epilogue warps (128 threads) =====
        const int q = warp & 3;                       // TMEM lane quarter this warp may access
        const uint32_t tq = tmem_base + (uint32_t(32 * q) << 16);
        const int e = threadIdx.x - 64;               // 0..127
        const int r64 = 16 * q + lane;                // row of an M = 64 tile held by this thread (valid for lane < 16)
        const bool r64_ok = lane < 16 && r64 < kMaxSlice;
        constexpr int kItems = NB * kQuads;           // (sample, feature quad) items of the transposed passes
        float v[NB];
        // This CTA has delivered its part of boundary i to every CTA of the cluster.  The release at cluster scope
        // by threads 0..7 covers the stores of the other epilogue threads (ordered before it by the CTA barrier);
        // the proxy fence makes the generic-proxy stores into shared memory visible to the tensor core's reads.
        auto deliver = [&](int i) {
            fence_proxy_async_all();
            epi_bar_sync();
            if (e < CL) mbar_arrive_remote(map_to_cluster_rank(smem_u32(&xbar[i]), uint32_t(e)));
        };
        // `val` = features [fbase + 4 fq, +4) of sample n: into EVERY CTA's B tile at `tile_byte_off` (when `spread`),
        // and into the global activation tensor for the weight-gradient kernels.
        auto scatter = [&](int tile_byte_off, int fbase, int nfeat, float* gdst, int n, int fq, float4 val, bool spread) {
            const int k = fbase + 4 * fq;
            if (spread) {
                const uint32_t local = smem_u32(smem + tile_byte_off) + tile_off<NB>(n, k);
#pragma unroll
                for (int d = 0; d < CL; d++) st_cluster_f32x4(map_to_cluster_rank(local, uint32_t(d)), val.x, val.y, val.z, val.w);
            }
            if (n0 + n < a.m) *reinterpret_cast<float4*>(gdst + size_t(n0 + n) * nfeat + k) = val;
        };
        auto quad = [&](const float* buf, int n, int fq) {  // features 4 fq .. 4 fq + 3 of sample n out of a [feature][sample] buffer
            return make_float4(buf[(4 * fq) * NB + n], buf[(4 * fq + 1) * NB + n], buf[(4 * fq + 2) * NB + n], buf[(4 * fq + 3) * NB + n]);
        };
        // this thread's M = 64 row: v[n] * (1 - h[row][n]^2) into the transposition buffer
        auto times_tanh_grad = [&](const float* hbuf) {
            if (!r64_ok) return;
#pragma unroll
            for (int i = 0; i < NB / 4; i++) {
                const float4 h = *reinterpret_cast<const float4*>(hbuf + r64 * NB + 4 * i);
                *reinterpret_cast<float4*>(stg + r64 * NB + 4 * i) =
                    make_float4(v[4 * i] * (1.f - h.x * h.x), v[4 * i + 1] * (1.f - h.y * h.y), v[4 * i + 2] * (1.f - h.z * h.z), v[4 * i + 3] * (1.f - h.w * h.w));
            }
        };

        if (!MID) {
        // ---- epilogue 1a: reduce-scatter the partial h1^T tiles (feature f goes to the CTA that owns slice f / s1) ----
            mbar_wait(&acc_bar[0], 0);
            if (e == 0) CHAIN_STAMP(13);
            tcgen05_fence_after();
            for (int t = 0; t < t1; t++) {
                __syncwarp();
                tmem_ld_row<NB>(tq + 32 * t, v);
                if (kbA == kbB) {  // a short K gave this CTA no K block: its partial is zero, not an unwritten accumulator
#pragma unroll
                    for (int i = 0; i < NB; i++) v[i] = 0.f;
                }
                const int f = 128 * t + 32 * q + lane;
                if (f < a.n1) {
                    const int d = f / s1, fl = f - d * s1;
                    const uint32_t dst = map_to_cluster_rank(smem_u32(red + (size_t(c) * kMaxSlice + fl) * NB), uint32_t(d));
#pragma unroll
                    for (int i = 0; i < NB / 4; i++) st_cluster_f32x4(dst + 16 * i, v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                }
            }
            epi_bar_sync();
            if (e < CL) mbar_arrive_remote(map_to_cluster_rank(smem_u32(&xbar[0]), uint32_t(e)));
            mbar_wait_cluster(&xbar[0], 0);
            if (e == 0) CHAIN_STAMP(14);
            // ---- epilogue 1b: h1 slice = tanh(sum over the 8 partials in CTA order + b1) -> every CTA's T0 ----
            for (int item = e; item < kItems; item += 128) {
                const int n = item / kQuads, fq = item % kQuads;
                const int k = f1 + 4 * fq;
                if (4 * fq >= s1 || k >= a.n1) continue;
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
                for (int src = 0; src < CL; src++)
#pragma unroll
                    for (int i = 0; i < 4; i++) acc[i] += red[(size_t(src) * kMaxSlice + 4 * fq + i) * NB + n];
                const float4 b = __ldg(reinterpret_cast<const float4*>(a.b1 + k));
                const float4 h = make_float4(tanhf(acc[0] + b.x), tanhf(acc[1] + b.y), tanhf(acc[2] + b.z), tanhf(acc[3] + b.w));
                h1s[(4 * fq + 0) * NB + n] = h.x;
                h1s[(4 * fq + 1) * NB + n] = h.y;
                h1s[(4 * fq + 2) * NB + n] = h.z;
                h1s[(4 * fq + 3) * NB + n] = h.w;
                scatter(kOffT0, f1, a.n1, a.h1, n, fq, h, stop >= 2);
            }
            if (stop >= 2) deliver(1);  // (its CTA barrier also orders everybody's reads of `red` = T1 before the peers may fill T1)
        } else {
            // MID: this CTA's slice of h1 (for 1 - h1^2 in epilogue 5) comes from global memory
            for (int item = e; item < kItems; item += 128) {
                const int n = item / kQuads, fq = item % kQuads;
                const int k = f1 + 4 * fq;
                if (4 * fq >= s1 || k >= a.n1) continue;
                const float4 h = (n0 + n < a.m) ? __ldg(reinterpret_cast<const float4*>(a.h1 + size_t(n0 + n) * a.n1 + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
                h1s[(4 * fq + 0) * NB + n] = h.x;
                h1s[(4 * fq + 1) * NB + n] = h.y;
                h1s[(4 * fq + 2) * NB + n] = h.z;
                h1s[(4 * fq + 3) * NB + n] = h.w;
            }
        }
        if (e == 0) CHAIN_STAMP(15);
        if (stop < 2) goto epi_done;
        // ---- epilogue 2: h2 slice = tanh(acc + b2) -> every CTA's T1 ----
        {
            const int g = f2 + r64;
            const bool ok = r64_ok && r64 < s2 && g < a.n2;
            const float bias = ok ? __ldg(a.b2 + g) : 0.f;
            mbar_wait(&acc_bar[1], 0);
            if (e == 0) CHAIN_STAMP(16);
            tcgen05_fence_after();
            tmem_ld_row<NB>(tq + 64, v);
            if (r64_ok) {
#pragma unroll
                for (int i = 0; i < NB / 4; i++)
                    *reinterpret_cast<float4*>(h2s + r64 * NB + 4 * i) =
                        ok ? make_float4(tanhf(v[4 * i] + bias), tanhf(v[4 * i + 1] + bias), tanhf(v[4 * i + 2] + bias), tanhf(v[4 * i + 3] + bias))
                           : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            epi_bar_sync();
            for (int item = e; item < kItems; item += 128) {
                const int n = item / kQuads, fq = item % kQuads;
                if (4 * fq < s2 && f2 + 4 * fq < a.n2) scatter(kOffT1, f2, a.n2, a.h2, n, fq, quad(h2s, n, fq), stop >= 3);
            }
            if (stop >= 3) deliver(2);
        }
        if (e == 0) CHAIN_STAMP(17);
        if (stop < 3) goto epi_done;
        // ---- epilogue 3: softmax + loss gradient + metrics.  The warp of lane quarter 0 moves the logits (classes on
        //      lanes) to shared memory; NB of its threads then own one sample each (the whole row in registers) ----
        if (q == 0) {
            mbar_wait(&acc_bar[2], 0);
            if (lane == 0) CHAIN_STAMP(18);
            tcgen05_fence_after();
            tmem_ld_row<NB>(tq + 96, v);
            if (lane < 16) {
#pragma unroll
                for (int n = 0; n < NB; n++) smx[lane * (NB + 1) + n] = v[n];
            }
            __syncwarp();
            float ce = 0.f, mse = 0.f, correct = 0.f, count = 0.f;
            if (lane < NB) {
                const int n = lane;
                const int label = (a.labels && n0 + n < a.m) ? int(__ldg(a.labels + n0 + n)) : -1;
                float p[16];
                float mx = -3.0e38f;
                int arg = 0;
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    p[i] = (i < a.nc) ? smx[i * (NB + 1) + n] + __ldg(a.b3 + i) : -3.0e38f;
                    if (mx < p[i]) {
                        mx = p[i];
                        arg = i;
                    }
                }
                float sum = 0.f;
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    p[i] = (i < a.nc) ? __expf(p[i] - mx) : 0.f;
                    sum += p[i];
                }
                const float inv = 1.0f / sum;
                float er[16];
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    p[i] *= inv;
                    er[i] = (i < a.nc) ? p[i] - (i == label ? 1.0f : 0.0f) : 0.0f;
                    mse += er[i] * er[i];
                    if (i == label) ce = -__logf(fmaxf(p[i], 1e-30f));
                }
                uint8_t* z3 = smem + kOffZ3;
#pragma unroll
                for (int i = 0; i < 4; i++)  // B operand tile [sample n][class], K-major, 128-byte swizzle
                    *reinterpret_cast<float4*>(z3 + n * 128 + ((i ^ (n & 7)) << 4)) = make_float4(er[4 * i], er[4 * i + 1], er[4 * i + 2], er[4 * i + 3]);
                if (c == 0 && n0 + n < a.m) {
#pragma unroll
                    for (int i = 0; i < 16; i++)
                        if (i < a.ldc) {
                            a.probs[size_t(n0 + n) * a.ldc + i] = p[i];
                            if (a.dz3) a.dz3[size_t(n0 + n) * a.ldc + i] = er[i];
                        }
                }
                correct = (label >= 0 && arg == label) ? 1.0f : 0.0f;
                count = label >= 0 ? 1.0f : 0.0f;
                if (label < 0) ce = mse = 0.f;
                mse /= float(a.nc);
            }
            ce = warp_sum(ce);
            mse = warp_sum(mse);
            correct = warp_sum(correct);
            count = warp_sum(count);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                CHAIN_STAMP(19);
                mbar_arrive(z3_bar);
                if (c == 0 && a.metrics && count > 0.f) {
                    atomicAdd(a.metrics + 0, ce);
                    atomicAdd(a.metrics + 1, mse);
                    atomicAdd(a.metrics + 2, correct);
                    atomicAdd(a.metrics + 3, count);
                }
            }
        }
        if (stop < 4) goto epi_done;
        // ---- epilogue 4: dz2 slice = acc * (1 - h2^2) -> every CTA's T0 ----
        mbar_wait(&acc_bar[3], 0);
        if (e == 0) CHAIN_STAMP(20);
        tcgen05_fence_after();
        tmem_ld_row<NB>(tq + 128, v);
        times_tanh_grad(h2s);
        epi_bar_sync();
        for (int item = e; item < kItems; item += 128) {
            const int n = item / kQuads, fq = item % kQuads;
            if (4 * fq < s2 && f2 + 4 * fq < a.n2) scatter(kOffT0, f2, a.n2, a.dz2, n, fq, quad(stg, n, fq), stop >= 5);
        }
        if (stop >= 5) deliver(3);
        if (e == 0) CHAIN_STAMP(21);
        if (stop < 5) goto epi_done;
        // ---- epilogue 5: dz1 slice = acc * (1 - h1^2) -> every CTA's T1 ----
        mbar_wait(&acc_bar[4], 0);
        if (e == 0) CHAIN_STAMP(22);
        tcgen05_fence_after();
        tmem_ld_row<NB>(tq + 160, v);
        // (everybody has read `stg` in epilogue 4's second pass: deliver(3)'s CTA barrier came after it)
        times_tanh_grad(h1s);
        epi_bar_sync();
        for (int item = e; item < kItems; item += 128) {
            const int n = item / kQuads, fq = item % kQuads;
            if (4 * fq < s1 && f1 + 4 * fq < a.n1) scatter(kOffT1, f1, a.n1, a.dz1, n, fq, quad(stg, n, fq), stop >= 6);
        }
        if (stop >= 6) deliver(4);
        if (e == 0) CHAIN_STAMP(23);
        if (stop < 6) goto epi_done;
        // ---- epilogue 6: dx = acc * act'(x), up to two M = 128 tiles (row = TMEM lane) ----
        {
            float xv[NB];
            for (int t = 0; t < xt; t++) {
                const int jr = 128 * t + 32 * q + lane, j = j0 + jr;
                const bool ok = jr < jn;
#pragma unroll
                for (int n = 0; n < NB; n++) xv[n] = (ok && n0 + n < a.m) ? __ldg(a.x + size_t(n0 + n) * a.k1 + j) : 0.f;
                if (t == 0) {
                    mbar_wait(&acc_bar[5], 0);
                    if (e == 0) CHAIN_STAMP(24);
                    tcgen05_fence_after();
                }
                __syncwarp();
                tmem_ld_row<NB>(tq + 192 + 32 * t, v);
                if (ok) {
#pragma unroll
                    for (int n = 0; n < NB; n++)
                        if (n0 + n < a.m) a.dx[size_t(n0 + n) * a.k1 + j] = v[n] * act_grad_from_output(a.act_x, xv[n]);
                }
            }
        }
    epi_done:;
        if (e == 0) CHAIN_STAMP(25);
    }
    tcgen05_fence_before();
    __syncthreads();
    cluster_sync_all();  // no CTA leaves while a peer may still write its shared memory or arrive on its barriers
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

// K-major operand map: tensor [rows][k] (k contiguous), box {32 k, box_rows}, 128-byte swizzle.
int kmajor_map(CUtensorMap* tm, const float* base, int rows, int k, int ld, int box_rows) {
    const uint64_t dims[2] = {uint64_t(k), uint64_t(rows)}, strides[2] = {0, uint64_t(ld) * 4};
    const uint32_t box[2] = {32, uint32_t(box_rows)};
    return encode_tmap_f32(tm, base, 2, dims, strides, box, nullptr, 128);
}
// M-major operand map over a row-major [k_rows][cols] tensor: box {32 cols, box_k rows}, 128-byte swizzle with 32-byte atoms.
int mmajor_map(CUtensorMap* tm, const float* base, int k_rows, int cols, int ld, int box_k) {
    const uint64_t dims[2] = {uint64_t(cols), uint64_t(k_rows)}, strides[2] = {0, uint64_t(ld) * 4};
    const uint32_t box[2] = {32, uint32_t(box_k)};
    return encode_tmap_f32(tm, base, 2, dims, strides, box, nullptr, kTmaSw128Atom32);
}

long long* g_chain_stamps = nullptr;  // b200cnn_fc_chain_debug_stamps()

}  // namespace

bool tc_fc_chain_supported(int m, int k1, int n1, int n2, int nc, int ldc) {
    return m > 0 && k1 >= 32 && k1 % 4 == 0 && k1 <= 8 * 224 && n1 >= 8 && n1 % 4 == 0 && n1 <= 208 && n2 >= 8 && n2 % 4 == 0 &&
           n2 <= 208 && nc >= 1 && nc <= 16 && ldc >= nc && ldc <= 16;
}

bool tc_fc_chain_enabled() {
    static const bool on = [] {
        // Opt-in (B200CNN_CHAIN=1).  Measured at batch 256 on B200: 33 us per launch (29 us inside the kernel: GEMM 1
        // 4.4, four layer boundaries ~1.5 each, 25-MMA GEMMs 1.6 each, GEMM 6 4, epilogues) against 38 us for the seven
        // kernels it replaces — but those overlap each other's prologues (programmatic dependent launch) and leave SMs
        // to the weight-gradient kernels on the side stream, which this 220 KB-per-CTA kernel cannot, so the whole
        // step goes from 71.4 to 84.5 us.  Kept as a tested building block; it pays off once a tcgen05.mma of this
        // size costs less than the ~65 ns measured here (A operand from TMEM) or batches grow.
        const char* v = getenv("B200CNN_CHAIN");
        return v && v[0] == '1';
    }();
    return on && current_backend() != B200CNN_BACKEND_SIMT;
}

int tc_fc_chain(const float* x, int act_x, const float* w1, const float* b1, const float* w2, const float* b2, const float* w3,
                const float* b3, const uint8_t* labels, float* h1, float* h2, float* probs, float* dz3, float* dz2, float* dz1,
                float* dx, float* metrics, int m, int k1, int n1, int n2, int nc, int ldc, cudaStream_t st) {
    if (!tc_fc_chain_supported(m, k1, n1, n2, nc, ldc)) return set_error(kInvalidValue, "tc_fc_chain: unsupported shape");
    for (const void* p : {(const void*)x, (const void*)w1, (const void*)w2, (const void*)w3, (const void*)h1, (const void*)h2,
                          (const void*)dz2, (const void*)dz1, (const void*)b1})
        if (reinterpret_cast<uintptr_t>(p) & 15) return set_error(kInvalidValue, "tc_fc_chain: operands must be 16-byte aligned");
    // One wave of clusters whenever possible: 16 samples per cluster up to 15 clusters, else 24 (see Cfg).
    const bool wide = ceil_div(m, 16) > 15;
    const int nb = wide ? 24 : 16;
    CUtensorMap tx, tw1a, tw1b, tw1m, tw2k, tw2m, tw3k, tw3m;
    int rc = 0;
    const int rows_b = n1 > 128 ? ((n1 - 128 + 7) & ~7) : 8;
    if ((rc = kmajor_map(&tx, x, m, k1, k1, nb)) || (rc = kmajor_map(&tw1a, w1, n1, k1, k1, 128)) ||
        (rc = kmajor_map(&tw1b, w1, n1, k1, k1, rows_b)) || (rc = mmajor_map(&tw1m, w1, n1, k1, k1, 32)) ||
        (rc = kmajor_map(&tw2k, w2, n2, n1, n1, 32)) || (rc = mmajor_map(&tw2m, w2, n2, n1, n1, 32)) ||
        (rc = kmajor_map(&tw3k, w3, nc, n2, n2, 16)) || (rc = mmajor_map(&tw3m, w3, nc, n2, n2, 16)))
        return rc;
    static const int stop = [] {
        const char* v = getenv("B200CNN_CHAIN_DBG");
        return v ? atoi(v) : 99;
    }();
    ChainArgs a{b1, b2, b3, labels, x, h1, h2, probs, dz3, dz2, dz1, dx, metrics, m, k1, n1, n2, nc, ldc, act_x, g_chain_stamps, stop};
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(tc_fc_chain_kernel<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg<16>::kSmem)));
        B200CNN_CUDA(cudaFuncSetAttribute(tc_fc_chain_kernel<24, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg<24>::kSmem)));
        configured = true;
    }
    const dim3 grid(1, ceil_div(m, nb), CL);
    if (wide)
        B200CNN_CUDA(launch_kernel_cluster(tc_fc_chain_kernel<24, false>, grid, dim3(kThreads), Cfg<24>::kSmem, st, CL, tx, tw1a, tw1b, tw1m, tw2k,
                                           tw2m, tw3k, tw3m, tx, a));
    else
        B200CNN_CUDA(launch_kernel_cluster(tc_fc_chain_kernel<16, false>, grid, dim3(kThreads), Cfg<16>::kSmem, st, CL, tx, tw1a, tw1b, tw1m, tw2k,
                                           tw2m, tw3k, tw3m, tx, a));
    B200CNN_CHECK_LAUNCH("tc_fc_chain");
    return 0;
}

// ---- the middle of the chain as its own kernel: h2, softmax / loss, dz3, dz2, dz1 from h1 ------------------------------
bool tc_fc_mid_supported(int m, int n1, int n2, int nc, int ldc) { return tc_fc_chain_supported(m, 32, n1, n2, nc, ldc); }

bool tc_fc_mid_enabled() {
    static const bool on = [] {
        // Opt-in (B200CNN_MID=1).  Correct (tests/test_gpu_ops.py::test_fc_mid_kernel) and three launches fewer, but the
        // step is SLOWER with it on B200: 80.1 vs 71.8 us.  The four kernels it replaces overlap their neighbours'
        // prologues and let the weight-gradient kernels of the side stream start one by one; the fused kernel releases
        // all of them at once at its end, where they compete with the tail of the critical path.
        const char* v = getenv("B200CNN_MID");
        return v && v[0] == '1';
    }();
    return on && current_backend() != B200CNN_BACKEND_SIMT;
}

int tc_fc_mid(const float* h1, const float* w2, const float* b2, const float* w3, const float* b3, const uint8_t* labels, float* h2,
              float* probs, float* dz3, float* dz2, float* dz1, float* metrics, int m, int n1, int n2, int nc, int ldc, cudaStream_t st) {
    if (!tc_fc_mid_supported(m, n1, n2, nc, ldc)) return set_error(kInvalidValue, "tc_fc_mid: unsupported shape");
    for (const void* p : {(const void*)h1, (const void*)w2, (const void*)w3, (const void*)h2, (const void*)dz2, (const void*)dz1})
        if (reinterpret_cast<uintptr_t>(p) & 15) return set_error(kInvalidValue, "tc_fc_mid: operands must be 16-byte aligned");
    constexpr int NB = 16;
    CUtensorMap th1, tw2k, tw2m, tw3k, tw3m;
    int rc = 0;
    if ((rc = kmajor_map(&th1, h1, m, n1, n1, NB)) || (rc = kmajor_map(&tw2k, w2, n2, n1, n1, 32)) || (rc = mmajor_map(&tw2m, w2, n2, n1, n1, 32)) ||
        (rc = kmajor_map(&tw3k, w3, nc, n2, n2, 16)) || (rc = mmajor_map(&tw3m, w3, nc, n2, n2, 16)))
        return rc;
    ChainArgs a{nullptr, b2, b3, labels, nullptr, const_cast<float*>(h1), h2, probs, dz3, dz2, dz1, nullptr, metrics, m, 32, n1, n2, nc, ldc,
                B200CNN_ACT_NONE, g_chain_stamps, 99};
    using C = Cfg<NB, true>;
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(tc_fc_chain_kernel<NB, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(C::kSmem)));
        configured = true;
    }
    const dim3 grid(1, ceil_div(m, NB), CL);
    // (the W1 / x maps of the full chain are unused here: the h1 map stands in for them)
    B200CNN_CUDA(launch_kernel_cluster(tc_fc_chain_kernel<NB, true>, grid, dim3(kThreads), C::kSmem, st, CL, th1, th1, th1, th1, tw2k, tw2m, tw3k,
                                       tw3m, th1, a));
    B200CNN_CHECK_LAUNCH("tc_fc_mid");
    return 0;
}

}  // namespace b200cnn

extern "C" int b200cnn_fc_mid(const float* h1, const float* w2, const float* b2, const float* w3, const float* b3, const uint8_t* labels,
                              float* h2, float* probs, float* dz3, float* dz2, float* dz1, float* metrics, int m, int n1, int n2, int nc,
                              int ldc, b200cnn_stream stream) {
    return b200cnn::tc_fc_mid(h1, w2, b2, w3, b3, labels, h2, probs, dz3, dz2, dz1, metrics, m, n1, n2, nc, ldc, b200cnn::as_stream(stream));
}

// How many clusters of the chain kernel the device can hold at once (cudaOccupancyMaxActiveClusters); a batch of more
// than 16 x this many samples runs in waves.  -1 on error.
extern "C" int b200cnn_fc_chain_max_active_clusters(void) {
    using namespace b200cnn;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(1, 64, CL);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = Cfg<16>::kSmem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = CL;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (cudaFuncSetAttribute(tc_fc_chain_kernel<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg<16>::kSmem)) != cudaSuccess) return -1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, tc_fc_chain_kernel<16, false>, &cfg) != cudaSuccess) return -1;
    return n;
}

// Debugging aid: device buffer of >= 32 int64 that CTA (0,0,0) of every later launch fills with SM-clock stamps of its
// milestones (NULL turns it off).  See tools/diag_chain.py.
extern "C" void b200cnn_fc_chain_debug_stamps(long long* device_buffer) { b200cnn::g_chain_stamps = device_buffer; }

extern "C" int b200cnn_fc_chain(const float* x, int act_x, const float* w1, const float* b1, const float* w2, const float* b2,
                                const float* w3, const float* b3, const uint8_t* labels, float* h1, float* h2, float* probs,
                                float* dz3, float* dz2, float* dz1, float* dx, float* metrics, int m, int k1, int n1, int n2, int nc,
                                int ldc, b200cnn_stream stream) {
    return b200cnn::tc_fc_chain(x, act_x, w1, b1, w2, b2, w3, b3, labels, h1, h2, probs, dz3, dz2, dz1, dx, metrics, m, k1, n1, n2, nc,
                                ldc, b200cnn::as_stream(stream));
}
