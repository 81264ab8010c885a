// The "head" of a training step in ONE kernel: for a tile of 128 samples
//     h2    = tanh(h1 W2^T + b2)                         (forward, hidden FC layer)
//     p     = softmax(h2 W3^T + b3);  dz3 = p - onehot   (forward output layer + loss gradient + metrics)
//     dz2   = (dz3 W3) * (1 - h2^2)                      (backward-data of the output layer)
//     dz1   = (dz2 W2) * (1 - h1^2)                      (backward-data of the hidden layer)
// Four chained tcgen05 GEMMs whose intermediate operands never leave the SM: each epilogue writes its
// result straight into shared memory in the swizzled K-major layout the next tcgen05.mma reads (plus one
// copy to global memory for the weight-gradient kernels).  Weights stream through a 3-stage TMA ring:
// W2 K-major (GEMM 1), W3 K-major (GEMM 2), W3 N-major (GEMM 3), W2 N-major (GEMM 4) — the same two
// row-major [out][in] tensors the reference stores (/root/reference/cnn.c:116-121), no transposed copies.
//
// Status: correct (tests/test_gpu_ops.py::test_fused_head_kernel) but NOT the default.  At batch 256 it runs on 2 CTAs and
// takes 45 us, the five separate kernels it replaces 26 us (they spread over 13-26 CTAs each): the per-thread row-strided
// epilogue traffic of 128 x 200 tiles concentrated on two SMs costs more than four kernel boundaries.  B200CNN_HEAD=1
// enables it in the engine; it pays off only when M-tiles >= SM count.
// Replaces, per mini-batch, Layer_feedForw_full x2 + softmax (cnn.c:113-152), the loss gradient
// (cnn.c:285-286) and the error-propagation half of Layer_feedBack_full x2 (cnn.c:158-166).
#include "backend.cuh"
#include "common.cuh"
#include "tc_common.cuh"

namespace b200cnn {

namespace {

using namespace tc;

constexpr int kThreads = 192;       // warp 0 TMA, warp 1 TMEM + MMA, warps 2-5 epilogue
constexpr int kRows = 128;          // samples per CTA (UMMA M)
constexpr int kMaxKb = 7;           // hidden sizes up to 7 x 32 = 224 (operand tiles), 208 (accumulator width)
constexpr int kTile = kRows * 128;  // one 32-float K block of an A operand: 16 KiB
constexpr int kStageBytes = 28 * 1024;
constexpr int kRing = 3;
constexpr int kAccOut = 208, kAcc3 = 224;  // TMEM columns: [0,208) acc1/acc4, [208,224) logits, [224,432) acc3

struct HeadArgs {
    const float* h1;   // [m][n1]  input activations (tanh outputs of the layer below)
    const float* b2;   // [n2]
    const float* b3;   // [nc]
    const uint8_t* labels;
    float* h2;         // [m][n2]
    float* probs;      // [m][ldc]
    float* dz3;        // [m][ldc]
    float* dz2;        // [m][n2]
    float* dz1;        // [m][n1]
    float* metrics;    // [4]
    int m, n1, n2, nc, ldc;
};

// byte offset of float column c of row r inside a K-major, 128-byte-swizzled A operand made of 16 KiB tiles
__device__ __forceinline__ uint32_t a_off(int r, int c) {
    return uint32_t((c >> 5) * kTile + r * 128 + ((((c & 31) >> 2) ^ (r & 7)) << 4) + (c & 3) * 4);
}

__global__ void __launch_bounds__(kThreads, 1)
tc_head_kernel(const __grid_constant__ CUtensorMap tm_h1, const __grid_constant__ CUtensorMap tm_w2k,
               const __grid_constant__ CUtensorMap tm_w3k, const __grid_constant__ CUtensorMap tm_w3n,
               const __grid_constant__ CUtensorMap tm_w2n, HeadArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_region = smem;                         // 7 tiles: h1, then h2, then dz2
    uint8_t* z3_tile = smem + kMaxKb * kTile;         // dz3 as a 1-tile A operand
    uint8_t* ring = z3_tile + kTile;                  // 3 x 28 KiB weight stages
    uint64_t* bars = reinterpret_cast<uint64_t*>(ring + kRing * kStageBytes);
    uint64_t* a_full = bars;            // h1 landed
    uint64_t* full = bars + 1;          // [kRing]
    uint64_t* empty = full + kRing;     // [kRing]
    uint64_t* acc_full = empty + kRing; // [4] accumulator of GEMM g complete
    uint64_t* op_ready = acc_full + 4;  // [3] epilogue g has written the next GEMM's A operand
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(op_ready + 3);

    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    const int m0 = blockIdx.x * kRows;
    const int kb1 = (a.n1 + 31) / 32, kb2 = (a.n2 + 31) / 32;  // K blocks of GEMM 1 / of GEMMs 2 and 4
    const int n1p = (a.n1 + 15) & ~15, n2p = (a.n2 + 15) & ~15;  // UMMA N (multiples of 16)
    const int g1 = (a.n1 + 31) / 32, g2 = (a.n2 + 31) / 32;      // 32-column groups of an N-major tile

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm_h1);
        prefetch_tmap(&tm_w2k);
        prefetch_tmap(&tm_w3k);
        prefetch_tmap(&tm_w3n);
        prefetch_tmap(&tm_w2n);
        mbar_init(a_full, 1);
        for (int s = 0; s < kRing; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int g = 0; g < 4; g++) mbar_init(&acc_full[g], 1);
        for (int g = 0; g < 3; g++) mbar_init(&op_ready[g], 128);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        // ===== TMA producer: h1, then 2 + kb1 + kb2 weight stages through the ring =====
        if (lane == 0) {
            mbar_arrive_expect_tx(a_full, uint32_t(kb1 * kTile));
            for (int kb = 0; kb < kb1; kb++) tma_load_2d(a_region + kb * kTile, &tm_h1, a_full, kb * 32, m0);
            int fill = 0;
            auto acquire = [&](uint32_t bytes) -> uint8_t* {
                const int s = fill % kRing;
                mbar_wait(&empty[s], ((fill / kRing) & 1) ^ 1);
                mbar_arrive_expect_tx(&full[s], bytes);
                return ring + s * kStageBytes;
            };
            for (int kb = 0; kb < kb1; kb++, fill++) {  // GEMM 1: W2 rows (out features) x 32 k, K-major
                uint8_t* dst = acquire(uint32_t(kAccOut * 128));
                tma_load_2d(dst, &tm_w2k, &full[fill % kRing], kb * 32, 0);
            }
            {  // GEMM 2: W3 [16 rows][kb2 x 32], K-major, all K blocks in one stage
                uint8_t* dst = acquire(uint32_t(kb2 * 16 * 128));
                for (int kb = 0; kb < kb2; kb++) tma_load_2d(dst + kb * 2048, &tm_w3k, &full[fill % kRing], kb * 32, 0);
                fill++;
            }
            {  // GEMM 3: W3 as N-major B: [32 k rows (classes)][g2 groups of 32 columns]
                uint8_t* dst = acquire(uint32_t(g2 * 4096));
                for (int g = 0; g < g2; g++) tma_load_2d(dst + g * 4096, &tm_w3n, &full[fill % kRing], g * 32, 0);
                fill++;
            }
            for (int kb = 0; kb < kb2; kb++, fill++) {  // GEMM 4: W2 as N-major B, one K block (32 rows of W2) per stage
                uint8_t* dst = acquire(uint32_t(g1 * 4096));
                for (int g = 0; g < g1; g++) tma_load_2d(dst + g * 4096, &tm_w2n, &full[fill % kRing], g * 32, kb * 32);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            const uint32_t a_addr = smem_u32(a_region), z_addr = smem_u32(z3_tile), r_addr = smem_u32(ring);
            int use = 0;
            auto stage_wait = [&]() -> uint32_t {
                const int s = use % kRing;
                mbar_wait(&full[s], (use / kRing) & 1);
                tcgen05_fence_after();
                return r_addr + s * kStageBytes;
            };
            auto stage_release = [&]() {
                umma_commit(&empty[use % kRing]);
                use++;
            };
            // GEMM 1: acc1 = h1 . W2^T
            mbar_wait(a_full, 0);
            tcgen05_fence_after();
            const uint32_t idesc1 = make_idesc_tf32(kRows, n2p, false, false);
            for (int kb = 0; kb < kb1; kb++) {
                const uint32_t b = stage_wait();
#pragma unroll
                for (int kk = 0; kk < 4; kk++)
                    umma_tf32(tmem_base, make_smem_desc(a_addr + kb * kTile + kk * 32, 16, 1024, kSw128),
                              make_smem_desc(b + kk * 32, 16, 1024, kSw128), idesc1, (kb | kk) != 0);
                stage_release();
            }
            umma_commit(&acc_full[0]);
            // GEMM 2: logits = h2 . W3^T   (A written by epilogue 1)
            mbar_wait(&op_ready[0], 0);
            tcgen05_fence_after();
            {
                const uint32_t idesc2 = make_idesc_tf32(kRows, 16, false, false);
                const uint32_t b = stage_wait();
                for (int kb = 0; kb < kb2; kb++)
#pragma unroll
                    for (int kk = 0; kk < 4; kk++)
                        umma_tf32(tmem_base + kAccOut, make_smem_desc(a_addr + kb * kTile + kk * 32, 16, 1024, kSw128),
                                  make_smem_desc(b + kb * 2048 + kk * 32, 16, 1024, kSw128), idesc2, (kb | kk) != 0);
                stage_release();
            }
            umma_commit(&acc_full[1]);
            // GEMM 3: acc3 = dz3 . W3   (K = 16 classes, B N-major)
            mbar_wait(&op_ready[1], 0);
            tcgen05_fence_after();
            {
                const uint32_t idesc3 = make_idesc_tf32(kRows, n2p, false, true);
                const uint32_t b = stage_wait();
#pragma unroll
                for (int kk = 0; kk < 2; kk++)
                    umma_tf32(tmem_base + kAcc3, make_smem_desc(z_addr + kk * 32, 16, 1024, kSw128),
                              make_smem_desc(b + kk * 1024, 4096, 512, kSw128Base32), idesc3, kk != 0);
                stage_release();
            }
            umma_commit(&acc_full[2]);
            // GEMM 4: acc4 = dz2 . W2   (A written in place by epilogue 3, B N-major)
            mbar_wait(&op_ready[2], 0);
            tcgen05_fence_after();
            const uint32_t idesc4 = make_idesc_tf32(kRows, n1p, false, true);
            for (int kb = 0; kb < kb2; kb++) {
                const uint32_t b = stage_wait();
#pragma unroll
                for (int kk = 0; kk < 4; kk++)
                    umma_tf32(tmem_base, make_smem_desc(a_addr + kb * kTile + kk * 32, 16, 1024, kSw128),
                              make_smem_desc(b + kk * 1024, 4096, 512, kSw128Base32), idesc4, (kb | kk) != 0);
                stage_release();
            }
            umma_commit(&acc_full[3]);
        }
    } else {
        // ===== epilogue warps: thread = one sample row =====
        const int q = warp & 3;
        const int r = 32 * q + lane;
        const int m = m0 + r;
        const bool live = m < a.m;
        const uint32_t lane_base = tmem_base + (uint32_t(32 * q) << 16);
        float v[16];

        // --- epilogue 1: h2 = tanh(acc1 + b2) -> global + A operand of GEMM 2 ---
        mbar_wait(&acc_full[0], 0);
        tcgen05_fence_after();
        for (int c0 = 0; c0 < kb2 * 32; c0 += 16) {
            __syncwarp();
            if (c0 < n2p) tmem_ld_32x32b_x16(lane_base + uint32_t(c0), v);
#pragma unroll
            for (int i = 0; i < 16; i++) {
                const int c = c0 + i;
                v[i] = (c < a.n2 && c0 < n2p) ? tanhf(v[i] + __ldg(a.b2 + c)) : 0.0f;  // zero K padding: 0 * w, never NaN * 0
            }
#pragma unroll
            for (int i = 0; i < 4; i++)
                *reinterpret_cast<float4*>(a_region + a_off(r, c0 + 4 * i)) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            if (live) {
                float* out = a.h2 + size_t(m) * a.n2 + c0;
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if (c0 + 4 * i + 3 < a.n2)
                        *reinterpret_cast<float4*>(out + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                    else
                        for (int j = 0; j < 4; j++)
                            if (c0 + 4 * i + j < a.n2) out[4 * i + j] = v[4 * i + j];
            }
        }
        fence_proxy_async_smem();
        tcgen05_fence_before();
        mbar_arrive(&op_ready[0]);

        // --- epilogue 2: softmax, loss gradient, metrics; dz3 -> A operand of GEMM 3 ---
        mbar_wait(&acc_full[1], 0);
        tcgen05_fence_after();
        __syncwarp();
        tmem_ld_32x32b_x16(lane_base + uint32_t(kAccOut), v);
        float ce = 0.f, mse = 0.f, correct = 0.f, count = 0.f;
        {
            float mx = -3.0e38f;
            int arg = 0;
#pragma unroll
            for (int i = 0; i < 16; i++)
                if (i < a.nc) {
                    v[i] += __ldg(a.b3 + i);
                    if (mx < v[i]) {
                        mx = v[i];
                        arg = i;
                    }
                }
            float sum = 0.f;
#pragma unroll
            for (int i = 0; i < 16; i++)
                if (i < a.nc) {
                    v[i] = __expf(v[i] - mx);
                    sum += v[i];
                }
            const float inv = 1.0f / sum;
            const int label = (live && a.labels) ? int(a.labels[m]) : -1;
            float e[16];
#pragma unroll
            for (int i = 0; i < 16; i++) {
                const float pr = i < a.nc ? v[i] * inv : 0.0f;
                e[i] = (i < a.nc && live) ? pr - (i == label ? 1.0f : 0.0f) : 0.0f;
                if (live && i < a.ldc) {
                    a.probs[size_t(m) * a.ldc + i] = pr;
                    a.dz3[size_t(m) * a.ldc + i] = e[i];
                }
                mse += e[i] * e[i];
                if (i == label) ce = -__logf(fmaxf(pr, 1e-30f));
            }
            if (live) {
                mse /= a.nc;
                correct = (arg == label) ? 1.0f : 0.0f;
                count = 1.0f;
            }
            // K-major tile [128 rows][32 floats]: columns 0..15 = dz3, 16..31 = 0
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const float4 val = i < 4 ? make_float4(e[4 * i], e[4 * i + 1], e[4 * i + 2], e[4 * i + 3]) : make_float4(0.f, 0.f, 0.f, 0.f);
                *reinterpret_cast<float4*>(z3_tile + r * 128 + ((i ^ (r & 7)) << 4)) = val;
            }
        }
        fence_proxy_async_smem();
        tcgen05_fence_before();
        mbar_arrive(&op_ready[1]);
        if (a.metrics) {
            ce = warp_sum(ce);
            mse = warp_sum(mse);
            correct = warp_sum(correct);
            count = warp_sum(count);
            if (lane == 0 && count > 0.f) {
                atomicAdd(a.metrics + 0, ce);
                atomicAdd(a.metrics + 1, mse);
                atomicAdd(a.metrics + 2, correct);
                atomicAdd(a.metrics + 3, count);
            }
        }

        // --- epilogue 3: dz2 = acc3 * (1 - h2^2), in place over h2 in the A region, + global ---
        mbar_wait(&acc_full[2], 0);
        tcgen05_fence_after();
        for (int c0 = 0; c0 < kb2 * 32; c0 += 16) {
            __syncwarp();
            if (c0 < n2p) tmem_ld_32x32b_x16(lane_base + uint32_t(kAcc3 + c0), v);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                float4* cell = reinterpret_cast<float4*>(a_region + a_off(r, c0 + 4 * i));
                const float4 h = *cell;  // h2 (0 in the K padding)
                float4 d;
                d.x = (c0 + 4 * i + 0 < a.n2 && c0 < n2p) ? v[4 * i + 0] * (1.0f - h.x * h.x) : 0.0f;
                d.y = (c0 + 4 * i + 1 < a.n2 && c0 < n2p) ? v[4 * i + 1] * (1.0f - h.y * h.y) : 0.0f;
                d.z = (c0 + 4 * i + 2 < a.n2 && c0 < n2p) ? v[4 * i + 2] * (1.0f - h.z * h.z) : 0.0f;
                d.w = (c0 + 4 * i + 3 < a.n2 && c0 < n2p) ? v[4 * i + 3] * (1.0f - h.w * h.w) : 0.0f;
                *cell = d;
                if (live) {
                    float* out = a.dz2 + size_t(m) * a.n2 + c0 + 4 * i;
                    if (c0 + 4 * i + 3 < a.n2) *reinterpret_cast<float4*>(out) = d;
                    else {
                        if (c0 + 4 * i + 0 < a.n2) out[0] = d.x;
                        if (c0 + 4 * i + 1 < a.n2) out[1] = d.y;
                        if (c0 + 4 * i + 2 < a.n2) out[2] = d.z;
                    }
                }
            }
        }
        fence_proxy_async_smem();
        tcgen05_fence_before();
        mbar_arrive(&op_ready[2]);

        // --- epilogue 4: dz1 = acc4 * (1 - h1^2) -> global ---
        mbar_wait(&acc_full[3], 0);
        tcgen05_fence_after();
        for (int c0 = 0; c0 < n1p; c0 += 16) {
            __syncwarp();
            tmem_ld_32x32b_x16(lane_base + uint32_t(c0), v);
            if (live) {
                const float* hrow = a.h1 + size_t(m) * a.n1 + c0;
                float* out = a.dz1 + size_t(m) * a.n1 + c0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    if (c0 + 4 * i + 3 < a.n1) {
                        const float4 h = __ldg(reinterpret_cast<const float4*>(hrow + 4 * i));
                        *reinterpret_cast<float4*>(out + 4 * i) =
                            make_float4(v[4 * i] * (1.f - h.x * h.x), v[4 * i + 1] * (1.f - h.y * h.y),
                                        v[4 * i + 2] * (1.f - h.z * h.z), v[4 * i + 3] * (1.f - h.w * h.w));
                    } else {
                        for (int j = 0; j < 4; j++)
                            if (c0 + 4 * i + j < a.n1) {
                                const float h = __ldg(hrow + 4 * i + j);
                                out[4 * i + j] = v[4 * i + j] * (1.f - h * h);
                            }
                    }
                }
            }
        }
        __syncwarp();
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

bool tc_head_supported(int m, int n1, int n2, int nc, int ldc) {
    return m > 0 && n1 >= 16 && n2 >= 16 && n1 <= kAccOut && n2 <= kAccOut && n1 % 4 == 0 && n2 % 4 == 0 && nc >= 1 && nc <= 16 &&
           ldc >= nc && ldc <= 16;
}

// h1 [m][n1] -> h2 [m][n2], probs / dz3 [m][ldc], dz2 [m][n2], dz1 [m][n1]; w2 [n2][n1], w3 [nc][n2].
int tc_head(const float* h1, const float* w2, const float* b2, const float* w3, const float* b3, const uint8_t* labels, float* h2,
            float* probs, float* dz3, float* dz2, float* dz1, float* metrics, int m, int n1, int n2, int nc, int ldc,
            cudaStream_t st) {
    if (!tc_head_supported(m, n1, n2, nc, ldc)) return set_error(kInvalidValue, "tc_head: unsupported shape");
    if (!aligned16(h1) || !aligned16(w2) || !aligned16(w3) || !aligned16(h2) || !aligned16(dz2) || !aligned16(dz1))
        return set_error(kInvalidValue, "tc_head: buffers must be 16-byte aligned");
    CUtensorMap t_h1, t_w2k, t_w3k, t_w3n, t_w2n;
    auto map2 = [](CUtensorMap* tm, const float* base, uint64_t d0, uint64_t d1, uint64_t ld, uint32_t b0, uint32_t b1, int sw) {
        const uint64_t dims[2] = {d0, d1}, strides[2] = {0, ld * 4};
        const uint32_t box[2] = {b0, b1};
        return encode_tmap_f32(tm, base, 2, dims, strides, box, nullptr, sw);
    };
    if (int rc = map2(&t_h1, h1, uint64_t(n1), uint64_t(m), uint64_t(n1), 32, kRows, 128)) return rc;          // A, K-major
    if (int rc = map2(&t_w2k, w2, uint64_t(n1), uint64_t(n2), uint64_t(n1), 32, kAccOut, 128)) return rc;      // B K-major: rows = out features
    if (int rc = map2(&t_w3k, w3, uint64_t(n2), uint64_t(nc), uint64_t(n2), 32, 16, 128)) return rc;           // B K-major: rows = classes
    if (int rc = map2(&t_w3n, w3, uint64_t(n2), uint64_t(nc), uint64_t(n2), 32, 32, kTmaSw128Atom32)) return rc;  // B N-major: k = classes
    if (int rc = map2(&t_w2n, w2, uint64_t(n1), uint64_t(n2), uint64_t(n1), 32, 32, kTmaSw128Atom32)) return rc;  // B N-major: k = out features
    HeadArgs a{h1, b2, b3, labels, h2, probs, dz3, dz2, dz1, metrics, m, n1, n2, nc, ldc};
    constexpr size_t smem = size_t(kMaxKb) * kTile + kTile + size_t(kRing) * kStageBytes + 1024 + 256;
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(tc_head_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        configured = true;
    }
    B200CNN_CUDA(launch_kernel(tc_head_kernel, dim3(ceil_div(m, kRows)), dim3(kThreads), smem, st, t_h1, t_w2k, t_w3k, t_w3n, t_w2n, a));
    return 0;
}

}  // namespace b200cnn
