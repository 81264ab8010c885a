// Fully connected layers on the 5th-generation tensor cores: one warp-specialised GEMM kernel
//   D[M][N] = sum_k A(m,k) * B(n,k)      TF32 operands (fp32 in memory), FP32 accumulation in TMEM
// fed by TMA (cp.async.bulk.tensor, 128-byte swizzle) through a 4-stage mbarrier ring, with the layer
// epilogues fused after tcgen05.ld:
//   forward          y  = act(x W^T + b)                     A = x   (K-major),  B = W  (K-major)
//   backward-data    dx = (dz W) * act'(a_prev)              A = dz  (K-major),  B = W  (N-major)
//   backward-weight  dW = dz^T x,  db = column sums of dz    A = dz  (M-major),  B = x  (N-major)
// The M-/N-major operand forms let all three read the SAME row-major tensors the reference uses
// ([batch][features] activations, [out][in] weights, /root/reference/cnn.c:116-121) — no transposed
// copies exist anywhere.  Replaces Layer_feedForw_full / Layer_feedBack_full (cnn.c:113-173).
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA issuer,
// warps 2-5 = epilogue (one TMEM lane quarter each; during backward-weight they also sum the staged
// dz tiles for the bias gradient).
#include "backend.cuh"
#include "common.cuh"
#include "tc_common.cuh"

namespace b200cnn {

namespace tc {

int encode_tmap_f32(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box, const uint32_t* elem_strides, int swizzle_bytes) {
    using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p)
            return set_error(kInvalidValue, "cuTensorMapEncodeTiled is not available from this driver");
        fn = reinterpret_cast<EncodeFn>(p);
    }
    const CUtensorMapSwizzle sw = swizzle_bytes == kTmaSw128Atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                  : swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                  : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                  : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                        : CU_TENSOR_MAP_SWIZZLE_NONE;
    cuuint64_t d[5], s[5];
    cuuint32_t b[5], e[5];
    for (int i = 0; i < rank; i++) {
        d[i] = dims[i];
        b[i] = box[i];
        e[i] = elem_strides ? elem_strides[i] : 1;
        if (i > 0) s[i - 1] = strides_bytes[i];
    }
    const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, cuuint32_t(rank), const_cast<void*>(base), d, s, b, e,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(kInvalidValue, "cuTensorMapEncodeTiled failed with CUresult %d", int(r));
    return 0;
}

}  // namespace tc

namespace {

using namespace tc;

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 32;   // floats: one 128-byte swizzle row
constexpr int UMMA_K = 8;     // TF32: 32 bytes of K per instruction
// Deeper rings for narrower tiles: at these sizes a CTA is bound by the latency of its own operand stream,
// so the ring is sized to keep ~150-190 KiB of TMA loads in flight.
template <int BLOCK_N>
constexpr int stages_for() { return BLOCK_N <= 16 ? 10 : BLOCK_N <= 32 ? 9 : BLOCK_N <= 64 ? 7 : 5; }
constexpr int kThreads = 192;
constexpr int kABytes = BLOCK_M * BLOCK_K * 4;  // 16 KiB per stage, either major

enum Epi { kEpiBiasAct = 0, kEpiActGrad = 1, kEpiStore = 2, kEpiSoftmax = 3 };

struct EpiArgs {
    float* C;
    int ldc;
    const float* bias;  // [N]                  (kEpiBiasAct)
    const float* aux;   // [M][ld_aux]          (kEpiActGrad)
    int ld_aux;
    int act;
    float* rowsum;      // [M]  (+)= sum_k A(m,k) (kEpiStore, A M-major)
    int accumulate;
    // kEpiSoftmax (N <= 16): C = probabilities, dz = probs - onehot(labels), metrics[4] += {CE, MSE, #correct, #rows}
    const uint8_t* labels;
    float* dz;
    float* metrics;
    int perm_p, perm_c; // kEpiActGrad: if perm_p > 0 store column n at (n % perm_p) * perm_c + n / perm_p
                        // (CHW-flattened -> HWC: hands the convolution below its NHWC gradient directly)
};

constexpr int kMaxSplitK = 4;
template <int BLOCK_N>
constexpr size_t smem_bytes() {
    return size_t(stages_for<BLOCK_N>()) * (kABytes + BLOCK_N * BLOCK_K * 4) + 1024 /*alignment slack*/ + 512 /*barriers*/ +
           (BLOCK_N == 16 ? size_t(kMaxSplitK - 1) * BLOCK_M * BLOCK_N * 4 : 0) /*split-K partials of the cluster peers*/;
}

template <int BLOCK_N, bool A_MN, bool B_MN, int EPI>
__global__ void __launch_bounds__(kThreads, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, EpiArgs ep, int M,
               int N, int K) {
    static_assert(BLOCK_N % 16 == 0 && BLOCK_N >= 16 && BLOCK_N <= 256, "UMMA N for M=128");
    static_assert(!B_MN || BLOCK_N % 32 == 0, "an N-major B tile is made of 32-float column groups");
    constexpr int kStages = stages_for<BLOCK_N>();
    constexpr int kBBytes = BLOCK_N * BLOCK_K * 4;
    constexpr int kStageBytes = kABytes + kBBytes;
    constexpr int kTmemCols = BLOCK_N < 32 ? 32 : BLOCK_N;  // power of two >= 32 (BLOCK_N is 16/32/64/128)
    constexpr bool kRowSum = (EPI == kEpiStore) && A_MN;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
    float* splitk_red = reinterpret_cast<float*>(smem + kStages * kStageBytes + 512);  // [peers][BLOCK_M][BLOCK_N]

    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    const int m0 = blockIdx.y * BLOCK_M, n0 = blockIdx.x * BLOCK_N;
    // Split-K: gridDim.z CTAs form one cluster; CTA z accumulates k-blocks [kb0, kb1) in its own TMEM, the
    // peers ship their partial tiles into CTA 0's shared memory (DSMEM), CTA 0 adds them and runs the epilogue.
    const int splits = int(gridDim.z), zrank = int(blockIdx.z);
    const int total_kb = (K + BLOCK_K - 1) / BLOCK_K;
    const int kb0 = total_kb * zrank / splits, kb1 = total_kb * (zrank + 1) / splits;
    const int num_kb = kb1 - kb0;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_a);
        prefetch_tmap(&tmap_b);
        for (int s = 0; s < kStages; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kRowSum ? 5 : 1);  // MMA commit (+ one arrival per epilogue warp)
        }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();               // everything above overlapped the previous kernel's tail
    pdl_launch_dependents();  // let the next kernel begin its own prologue
    float rs = 0.0f;          // bias-gradient partial of an epilogue thread (backward-weight only)
    int label_pre = -1;       // prefetched label of this thread's row (softmax epilogue only)
    // epilogue operands prefetched while the mainloop runs: activation values (backward-data) or the bias (forward)
    float auxr[((EPI == kEpiActGrad || EPI == kEpiBiasAct || EPI == kEpiSoftmax) && BLOCK_N <= 64) ? BLOCK_N : 1];

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            for (int kb = 0; kb < num_kb; kb++) {
                const int s = kb % kStages;
                const uint32_t ph = (kb / kStages) & 1;
                mbar_wait(&empty_bar[s], ph ^ 1);
                uint8_t* a_dst = smem + s * kStageBytes;
                uint8_t* b_dst = a_dst + kABytes;
                mbar_arrive_expect_tx(&full_bar[s], kStageBytes);
                const int k0 = (kb0 + kb) * BLOCK_K;
                if (A_MN) {
#pragma unroll
                    for (int g = 0; g < BLOCK_M / 32; g++) tma_load_2d(a_dst + g * 4096, &tmap_a, &full_bar[s], m0 + 32 * g, k0);
                } else {
                    tma_load_2d(a_dst, &tmap_a, &full_bar[s], k0, m0);
                }
                if (B_MN) {
#pragma unroll
                    for (int g = 0; g < BLOCK_N / 32; g++) tma_load_2d(b_dst + g * 4096, &tmap_b, &full_bar[s], n0 + 32 * g, k0);
                } else {
                    tma_load_2d(b_dst, &tmap_b, &full_bar[s], k0, n0);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(BLOCK_M, BLOCK_N, A_MN, B_MN);
            for (int kb = 0; kb < num_kb; kb++) {
                const int s = kb % kStages;
                const uint32_t ph = (kb / kStages) & 1;
                mbar_wait(&full_bar[s], ph);
                tcgen05_fence_after();
                const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
                const uint32_t b_addr = a_addr + kABytes;
#pragma unroll
                for (int kk = 0; kk < BLOCK_K / UMMA_K; kk++) {
                    // K-major: 8 floats further along the 128-byte row.  M/N-major: the next 8-row group.
                    // (TF32 M/N-major operands exist only in the 128B-swizzle-with-32B-atoms layout: 4-row
                    //  swizzle period, so the stride between k groups is 512 B.)
                    const uint64_t da = A_MN ? make_smem_desc(a_addr + kk * 1024, 4096, 512, kSw128Base32)
                                             : make_smem_desc(a_addr + kk * 32, 16, 1024, kSw128);
                    const uint64_t db = B_MN ? make_smem_desc(b_addr + kk * 1024, 4096, 512, kSw128Base32)
                                             : make_smem_desc(b_addr + kk * 32, 16, 1024, kSw128);
                    umma_tf32(tmem_base, da, db, idesc, (kb | kk) != 0);
                }
                umma_commit(&empty_bar[s]);  // frees the stage once these MMAs have read it
            }
            umma_commit(tmem_full_bar);
        }
    } else {
        // ===== epilogue warps =====
        const int q = warp & 3;             // TMEM lane quarter this warp may access
        const int row_in_tile = 32 * q + lane;
        if (kRowSum) {
            // Bias gradient: column sums of dz == row sums of the staged M-major A tile.  Row r of the
            // tile is one batch sample; this thread owns column `row_in_tile` (undoing the 32-byte-atom swizzle).
            const int g = row_in_tile / 32, j = row_in_tile % 32;
            for (int kb = 0; kb < num_kb; kb++) {
                const int s = kb % kStages;
                const uint32_t ph = (kb / kStages) & 1;
                mbar_wait(&full_bar[s], ph);
                if (blockIdx.x == 0) {
                    const uint8_t* a_tile = smem + s * kStageBytes + g * 4096;
#pragma unroll 8
                    for (int r = 0; r < BLOCK_K; r++)
                        rs += *reinterpret_cast<const float*>(a_tile + r * 128 + (((j >> 3) ^ (r & 3)) << 5) + (j & 7) * 4);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty_bar[s]);
            }
        }
        // backward-data: fetch this thread's slice of the activation values (for act') NOW, while TMA and the tensor
        // core are busy, instead of exposing an L2 round trip per 16-column chunk after the accumulator is ready.
        constexpr bool kPrefetchAux = (EPI == kEpiActGrad) && BLOCK_N <= 64;
        if (kPrefetchAux && ep.aux && m0 + row_in_tile < M) {
            const float* arow = ep.aux + size_t(m0 + row_in_tile) * ep.ld_aux + n0;
#pragma unroll
            for (int i = 0; i < (kPrefetchAux ? BLOCK_N : 1); i++) auxr[i] = (n0 + i < N) ? __ldg(arow + i) : 0.0f;
        }
        constexpr bool kPrefetchBias = (EPI == kEpiBiasAct || EPI == kEpiSoftmax) && BLOCK_N <= 64;
        if (kPrefetchBias) {
#pragma unroll
            for (int i = 0; i < (kPrefetchBias ? BLOCK_N : 1); i++) auxr[i] = (ep.bias && n0 + i < N) ? __ldg(ep.bias + n0 + i) : 0.0f;
        }
        if (EPI == kEpiSoftmax && ep.labels && m0 + row_in_tile < M) label_pre = int(ep.labels[m0 + row_in_tile]);
        mbar_wait(tmem_full_bar, 0);
        tcgen05_fence_after();
        if (BLOCK_N == 16 && splits > 1) {
            if (zrank != 0) {
                float v[16];
                tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16), v);
                const uint32_t dst = map_to_cluster_rank(smem_u32(splitk_red + (size_t(zrank - 1) * BLOCK_M + row_in_tile) * BLOCK_N), 0);
#pragma unroll
                for (int i = 0; i < 4; i++) st_cluster_f32x4(dst + 16 * i, v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            }
        }
    }
    if (BLOCK_N == 16 && splits > 1) {
        __syncwarp();
        cluster_sync_all();  // peers' partial tiles have landed in CTA 0
    }
    if (warp >= 2 && zrank == 0) {
        const int q = warp & 3;
        const int row_in_tile = 32 * q + lane;
        const int m = m0 + row_in_tile;
        if (EPI == kEpiSoftmax) {
            // Whole row (<= 16 logits) in this thread: softmax, cross-entropy gradient and the reference's
            // metrics (cnn.c:125-143, :275-286, :508-513) without another kernel or another pass over memory.
            float v[16];
            tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16), v);
            float ce = 0.f, mse = 0.f, correct = 0.f, count = 0.f;
            if (m < M) {
                float mx = -3.0e38f;
                int arg = 0;
#pragma unroll
                for (int i = 0; i < 16; i++)
                    if (i < N) {
                        v[i] += auxr[i];
                        if (mx < v[i]) {
                            mx = v[i];
                            arg = i;
                        }
                    }
                float sum = 0.f;
#pragma unroll
                for (int i = 0; i < 16; i++)
                    if (i < N) {
                        v[i] = __expf(v[i] - mx);
                        sum += v[i];
                    }
                const float inv = 1.0f / sum;
                const int label = label_pre;
                float* prow = ep.C + size_t(m) * ep.ldc;
                float* drow = ep.dz ? ep.dz + size_t(m) * ep.ldc : nullptr;
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    const float pr = i < N ? v[i] * inv : 0.0f;
                    const float e = i < N ? pr - (i == label ? 1.0f : 0.0f) : 0.0f;
                    if (i < ep.ldc) {
                        prow[i] = pr;
                        if (drow) drow[i] = e;
                    }
                    mse += e * e;
                    if (i == label) ce = -__logf(fmaxf(pr, 1e-30f));
                }
                mse /= N;
                correct = (arg == label) ? 1.0f : 0.0f;
                count = 1.0f;
            }
            if (ep.metrics && ep.labels) {
                ce = warp_sum(ce);
                mse = warp_sum(mse);
                correct = warp_sum(correct);
                count = warp_sum(count);
                if (lane == 0 && count > 0.f) {
                    atomicAdd(ep.metrics + 0, ce);
                    atomicAdd(ep.metrics + 1, mse);
                    atomicAdd(ep.metrics + 2, correct);
                    atomicAdd(ep.metrics + 3, count);
                }
            }
        } else
#pragma unroll(BLOCK_N <= 64 ? BLOCK_N / 16 : 1)
        for (int c0 = 0; c0 < BLOCK_N; c0 += 16) {
            float v[16];
            __syncwarp();  // tcgen05.ld is warp-collective: reconverge after the guarded stores below
            tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(c0), v);
            if (BLOCK_N == 16 && splits > 1) {
                for (int pz = 0; pz < splits - 1; pz++) {
                    const float4* pr = reinterpret_cast<const float4*>(splitk_red + (size_t(pz) * BLOCK_M + row_in_tile) * BLOCK_N);
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        const float4 t4 = pr[i];
                        v[4 * i] += t4.x; v[4 * i + 1] += t4.y; v[4 * i + 2] += t4.z; v[4 * i + 3] += t4.w;
                    }
                }
            }
            const int n_first = n0 + c0;
            if (m < M && n_first < N) {
                float* crow = ep.C + size_t(m) * ep.ldc + n_first;
                const int valid = min(16, N - n_first);
                if (EPI == kEpiBiasAct) {
#pragma unroll
                    for (int i = 0; i < 16; i++)
                        if (i < valid)
                            v[i] = apply_act(ep.act, v[i] + (BLOCK_N <= 64 ? auxr[(c0 + i) % (sizeof(auxr) / sizeof(float))]
                                                                           : (ep.bias ? __ldg(ep.bias + n_first + i) : 0.0f)));
                } else if (EPI == kEpiActGrad) {
                    if (ep.aux) {
                        if (EPI == kEpiActGrad && BLOCK_N <= 64) {
#pragma unroll
                            for (int i = 0; i < 16; i++)
                                if (i < valid) v[i] *= act_grad_from_output(ep.act, auxr[(c0 + i) % (sizeof(auxr) / sizeof(float))]);
                        } else {
                            const float* arow = ep.aux + size_t(m) * ep.ld_aux + n_first;
#pragma unroll
                            for (int i = 0; i < 16; i++)
                                if (i < valid) v[i] *= act_grad_from_output(ep.act, __ldg(arow + i));
                        }
                    }
                } else if (ep.accumulate) {
#pragma unroll
                    for (int i = 0; i < 16; i++)
                        if (i < valid) v[i] += crow[i];
                }
                if (EPI == kEpiActGrad && ep.perm_p > 0) {
                    float* orow = ep.C + size_t(m) * ep.ldc;
#pragma unroll
                    for (int i = 0; i < 16; i++)
                        if (i < valid) {
                            const int n = n_first + i;
                            orow[(n % ep.perm_p) * ep.perm_c + n / ep.perm_p] = v[i];
                        }
                } else if (valid == 16 && ((reinterpret_cast<uintptr_t>(crow) & 15) == 0)) {
#pragma unroll
                    for (int i = 0; i < 4; i++)
                        reinterpret_cast<float4*>(crow)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                } else {
#pragma unroll
                    for (int i = 0; i < 16; i++)
                        if (i < valid) crow[i] = v[i];
                }
            }
        }
        __syncwarp();
        if (kRowSum && blockIdx.x == 0 && ep.rowsum && m < M) ep.rowsum[m] = ep.accumulate ? ep.rowsum[m] + rs : rs;
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

// 2-D fp32 operand map.  K-major operand: tensor [rows][k] (k contiguous), box {32 k, tile_rows}.
// M/N-major operand: tensor [k][cols] (cols contiguous), box {32 cols, 32 k}.
int make_operand_map(CUtensorMap* tm, const float* base, bool mn_major, int rows_or_cols, int k_extent, int ld_floats,
                     int tile_rows) {
    uint64_t dims[2], strides[2] = {0, uint64_t(ld_floats) * 4};
    uint32_t box[2];
    if (mn_major) {
        dims[0] = uint64_t(rows_or_cols);
        dims[1] = uint64_t(k_extent);
        box[0] = 32;
        box[1] = BLOCK_K;
    } else {
        dims[0] = uint64_t(k_extent);
        dims[1] = uint64_t(rows_or_cols);
        box[0] = BLOCK_K;
        box[1] = uint32_t(tile_rows);
    }
    return encode_tmap_f32(tm, base, 2, dims, strides, box, nullptr, mn_major ? kTmaSw128Atom32 : 128);
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int BLOCK_N, bool A_MN, bool B_MN, int EPI>
int launch_cfg(const CUtensorMap& ta, const CUtensorMap& tb, const EpiArgs& ep, int M, int N, int K, cudaStream_t st,
               int splits = 1) {
    auto kern = tc_gemm_kernel<BLOCK_N, A_MN, B_MN, EPI>;
    constexpr size_t smem = smem_bytes<BLOCK_N>();
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        configured = true;
    }
    if (BLOCK_N != 16) splits = 1;
    const dim3 grid(ceil_div(N, BLOCK_N), ceil_div(M, BLOCK_M), splits);
    B200CNN_CUDA(launch_kernel_cluster(kern, grid, dim3(kThreads), smem, st, splits, ta, tb, ep, M, N, K));
    return 0;
}

// Tile width: the widest N tile that still yields >= 96 CTAs, because at these sizes a CTA's time is
// its own operand stream and more, narrower CTAs finish sooner.
int pick_block_n(int M, int N, int min_bn) {
    const int mt = ceil_div(M, BLOCK_M);
    for (int bn : {128, 64, 32, 16}) {
        if (bn < min_bn) break;
        if (mt * ceil_div(N, bn) >= 96) return bn;
    }
    return min_bn;
}

template <bool A_MN, bool B_MN, int EPI>
int dispatch(const float* A, int a_rows, int lda, const float* B, int b_rows, int ldb, const EpiArgs& ep, int M, int N, int K,
             cudaStream_t st) {
    (void)a_rows; (void)b_rows;
    const int bn = pick_block_n(M, N, B_MN ? 32 : 16);
    CUtensorMap ta, tb;
    if (int rc = make_operand_map(&ta, A, A_MN, M, K, lda, BLOCK_M)) return rc;
    if (int rc = make_operand_map(&tb, B, B_MN, N, K, ldb, bn)) return rc;
    switch (bn) {
        case 128: return launch_cfg<128, A_MN, B_MN, EPI>(ta, tb, ep, M, N, K, st);
        case 64: return launch_cfg<64, A_MN, B_MN, EPI>(ta, tb, ep, M, N, K, st);
        case 32: return launch_cfg<32, A_MN, B_MN, EPI>(ta, tb, ep, M, N, K, st);
        default:
            if constexpr (!B_MN) {
                // Long-K, few-CTA forward GEMMs (fc1: K = 1568 on 26 CTAs): 4-way split-K across a cluster.
                const int total_kb = ceil_div(K, BLOCK_K);
                const int ctas = ceil_div(N, 16) * ceil_div(M, BLOCK_M);
                const int splits = (EPI == kEpiBiasAct && total_kb >= 16 && ctas * kMaxSplitK <= 148) ? kMaxSplitK : 1;
                return launch_cfg<16, A_MN, B_MN, EPI>(ta, tb, ep, M, N, K, st, splits);
            }
            return set_error(kInvalidValue, "tc_gemm: no tile configuration");
    }
}

}  // namespace

// ---- shape predicates: TMA needs 16-byte aligned bases and row strides ------------------------------
bool tc_linear_fwd_supported(int m, int n, int k, int ldy) { return m > 0 && n > 0 && k > 0 && k % 4 == 0 && ldy >= n; }
bool tc_linear_bwd_data_supported(int m, int n, int k, int lddz) { return m > 0 && n > 0 && k > 0 && k % 4 == 0 && lddz % 4 == 0 && lddz >= n; }
bool tc_linear_bwd_weight_supported(int m, int n, int k, int lddz) { return m > 0 && n > 0 && k > 0 && k % 4 == 0 && lddz % 4 == 0 && lddz >= n; }

int tc_linear_fwd(const float* x, const float* w, const float* b, float* y, int m, int n, int k, int ldy, int act,
                  cudaStream_t st) {
    if (!aligned16(x) || !aligned16(w)) return set_error(kInvalidValue, "tc_linear_fwd: operands must be 16-byte aligned");
    EpiArgs ep{};
    ep.C = y; ep.ldc = ldy; ep.bias = b; ep.act = act;
    return dispatch<false, false, kEpiBiasAct>(x, m, k, w, n, k, ep, m, n, k, st);
}

bool tc_linear_softmax_supported(int m, int n, int k, int ld) { return m > 0 && n > 0 && n <= 16 && ld >= n && ld <= 16 && k > 0 && k % 4 == 0; }

int tc_linear_softmax_xent(const float* x, const float* w, const float* b, const uint8_t* labels, float* probs, float* dz,
                           float* metrics, int m, int n, int k, int ld, cudaStream_t st) {
    if (!aligned16(x) || !aligned16(w)) return set_error(kInvalidValue, "tc_linear_softmax_xent: operands must be 16-byte aligned");
    EpiArgs ep{};
    ep.C = probs; ep.ldc = ld; ep.bias = b; ep.labels = labels; ep.dz = dz; ep.metrics = metrics;
    CUtensorMap ta, tb;
    if (int rc = make_operand_map(&ta, x, false, m, k, k, BLOCK_M)) return rc;
    if (int rc = make_operand_map(&tb, w, false, n, k, k, 16)) return rc;
    return launch_cfg<16, false, false, kEpiSoftmax>(ta, tb, ep, m, n, k, st);
}

int tc_linear_bwd_data(const float* dz, const float* w, const float* a_prev, int act_prev, float* dx, int m, int n, int k,
                       int lddz, cudaStream_t st) {
    return tc_linear_bwd_data_permuted(dz, w, a_prev, act_prev, dx, m, n, k, lddz, 0, 0, st);
}

int tc_linear_bwd_data_permuted(const float* dz, const float* w, const float* a_prev, int act_prev, float* dx, int m, int n,
                                int k, int lddz, int perm_p, int perm_c, cudaStream_t st) {
    if (!aligned16(dz) || !aligned16(w)) return set_error(kInvalidValue, "tc_linear_bwd_data: operands must be 16-byte aligned");
    if (perm_p > 0 && perm_p * perm_c != k) return set_error(kInvalidValue, "tc_linear_bwd_data: permutation %dx%d does not cover %d columns", perm_p, perm_c, k);
    EpiArgs ep{};
    ep.C = dx; ep.ldc = k; ep.aux = a_prev; ep.ld_aux = k; ep.act = act_prev; ep.perm_p = perm_p; ep.perm_c = perm_c;
    // D[m][k_in] = sum over n_out: A = dz [m][n_out] K-major; B(n = k_in, k = n_out) = w[n_out][k_in] N-major.
    return dispatch<false, true, kEpiActGrad>(dz, m, lddz, w, k, k, ep, m, k, n, st);
}

int tc_linear_bwd_weight(const float* dz, const float* x, float* dw, float* db, int accumulate, int m, int n, int k,
                         int lddz, cudaStream_t st) {
    if (!aligned16(dz) || !aligned16(x)) return set_error(kInvalidValue, "tc_linear_bwd_weight: operands must be 16-byte aligned");
    EpiArgs ep{};
    ep.C = dw; ep.ldc = k; ep.rowsum = db; ep.accumulate = accumulate;
    // D[n_out][k_in] = sum over batch: A(m = n_out, k = batch) = dz[batch][n_out] M-major; B = x[batch][k_in] N-major.
    return dispatch<true, true, kEpiStore>(dz, n, lddz, x, k, k, ep, n, k, m, st);
}

}  // namespace b200cnn
