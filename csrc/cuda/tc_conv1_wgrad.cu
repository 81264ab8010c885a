// Weight gradient of the FIRST convolution (1 -> 16 channels, 3x3 / stride 2 / pad 1, 28x28 uint8 image -> 14x14)
// on the tensor cores.  Default path of the engine's tensor-core chain (B200CNN_TC_C1=0 falls back to the CUDA-core
// kernel in conv_nhwc.cu): 5.7 us per launch against 8.2 us for that kernel (tools/diag_c1_tc.py), the whole GPU
// test-suite passes with it, direct value test in tests/test_gpu_ops.py.  Its index algebra is pinned by
// tests/test_c1_tc_layout.py.
//
//     dW[oc][tap] = sum over (n, pixel) of dz[n][pixel][oc] * x[n][patch(pixel, tap)] / 255,   db[oc] = sum dz
//
// The reduction runs over 50,176 pixels per batch of 256 and the outputs are 16 x 9 numbers, so M and N of the
// GEMM have to be manufactured.  dz is NHWC with 16 channels: two horizontally adjacent pixels are one 128-byte
// row, i.e. [pixel pairs q][(parity e, oc)] is an M-major A operand that TMA loads as it is.  The matching B operand
// [pixel pairs q][(parity e', column j)] holds, for pixel 2q+e', its nine input taps (j = 0..8), the constant 1
// (j = 9: its product with dz is the bias gradient) and zeros (j = 10..15); the CTA's warps build it from the uint8
// image while the data-gradient kernel that produces dz is still running (it does not depend on dz).  Then
//     D[(e, oc)][(e', j)] = sum_q A[q][(e, oc)] * B[q][(e', j)],        dW[oc][j] = D[(0,oc)][(0,j)] + D[(1,oc)][(1,j)]
// (the e != e' blocks are by-products).  2 images per CTA = 196 pair rows (+4 zero rows) = 25 MMAs of M=64 (rows
// 32..63 unused), N=32, K=8.  The per-CTA result goes through the same "last block adds the partials in block order
// and applies the fused SGD step" tail as the CUDA-core kernel, so the result is reproducible run to run.
//
// Replaces the weight half of Layer_feedBack_conv (/root/reference/cnn.c:236,242) for the first convolution.
#include <cstdlib>

#include "../comm/xchg_device.cuh"
#include "backend.cuh"
#include "common.cuh"
#include "tc_common.cuh"

namespace b200cnn {

namespace {

using namespace tc;

constexpr int kThreads = 640;              // warp 0: TMA, warp 1: TMEM + MMA, warps 2..19: B tile; all: reduction tail
constexpr int kImgs = 2;                   // images per CTA
constexpr int kPairs = 98;                 // pixel pairs per 14x14 image
constexpr int kRows = kImgs * kPairs;      // 196 K rows
constexpr int kSteps = (kRows + 7) / 8;    // 25 (rows 196..199 are zero)
constexpr int kTile = kSteps * 8 * 128;    // 25,600 bytes
constexpr int kOC = 16, kTaps = 9, kNW = kOC * kTaps, kNOut = kNW + kOC;  // 160 results
constexpr int kInPix = 28 * 28;
constexpr int kBuilders = kThreads - 64;   // 576 threads build the B tile
constexpr int kParts = kThreads / kNOut;   // 4 groups of threads share the final sum
constexpr size_t kSmem = 2 * size_t(kTile) + kImgs * kInPix + 2 * kOC * 10 * 4 + kParts * kNOut * 4 + 64 + 1024;

struct C1Args {
    const uint8_t* x;   // [n][28][28]
    float* partials;    // [gridDim.x][160]
    float* dw;          // [16][9]
    float* db;          // [16]
    unsigned* counter;  // zero before the first launch; wraps back to zero
    float* sgd_w;       // optional fused update: w -= lr * dW, b -= lr * db
    float* sgd_b;
    const float* lr_dev;
    int n, accumulate;
    // Data parallel (xc.world > 1): the block that finishes last also all-reduces these 160 numbers with the peers and
    // applies the averaged SGD step to this rank's replica (one-shot: every rank adds all ranks' values in rank
    // order, so replicas stay bitwise identical).  dw / db then point into this rank's peer-visible gradient buffer,
    // w_off / b_off are their float offsets inside it (and of the parameters inside the parameter buffer).
    TailXchg xc;
    long w_off, b_off;
};

__global__ void __launch_bounds__(kThreads, 1)
tc_conv1_bwd_weight_kernel(const __grid_constant__ CUtensorMap tmap_dz, C1Args a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_tile = smem;            // dz pairs; the M=64 shape also reads one more group kTile further: the B tile
    uint8_t* b_tile = smem + kTile;    // taps
    uint8_t* xs = smem + 2 * kTile;    // the CTA's images
    float* part = reinterpret_cast<float*>(xs + kImgs * kInPix);  // [e][oc][10]
    float* red = part + 2 * kOC * 10;                             // [kParts][160]
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(red + kParts * kNOut);
    uint64_t* tmem_full_bar = full_bar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(full_bar + 2);
    __shared__ unsigned is_last;

    const int t = threadIdx.x, warp = t / 32, lane = t % 32;
    const int n0 = blockIdx.x * kImgs;
    const int ns = min(kImgs, a.n - n0);

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_dz);
        mbar_init(full_bar, 1);
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<32>(tmem_slot);
    if (warp >= 2) {
        // ---- B tile, independent of the producer of dz: runs before griddepcontrol.wait ----
        const int tb = t - 64;
        for (int i = tb; i < ns * kInPix / 4; i += kBuilders)
            reinterpret_cast<uchar4*>(xs)[i] = __ldg(reinterpret_cast<const uchar4*>(a.x + size_t(n0) * kInPix) + i);
        asm volatile("bar.sync 1, %0;" ::"n"(kBuilders) : "memory");
        // lane = column (e', j) of the 128-byte row: 32 lanes hit 32 different banks (32-byte chunk ^= row & 3)
        const int e = lane >> 4, j = lane & 15;
        const int ky = j / 3, kx = j % 3;
        for (int r = warp - 2; r < kRows + 4; r += kBuilders / 32) {
            float v = 0.0f;
            const int g = r / kPairs, q = r % kPairs;
            if (r < kRows && g < ns) {
                if (j < kTaps) {
                    const int y = q / 7, px = 2 * (q % 7) + e;
                    const int iy = 2 * y - 1 + ky, ix = 2 * px - 1 + kx;
                    if (iy >= 0 && iy < 28 && ix >= 0 && ix < 28) v = float(xs[g * kInPix + iy * 28 + ix]) * (1.0f / 255.0f);
                } else if (j == kTaps) {
                    v = 1.0f;
                }
            }
            *reinterpret_cast<float*>(b_tile + r * 128 + ((((lane >> 3) ^ (r & 3))) << 5) + (lane & 7) * 4) = v;
            if (r >= kRows)  // the four K rows TMA never writes
                *reinterpret_cast<float*>(a_tile + r * 128 + lane * 4) = 0.0f;
        }
        fence_proxy_async_smem();
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0) {
            mbar_arrive_expect_tx(full_bar, uint32_t(kRows * 128));
            tma_load_2d(a_tile, &tmap_dz, full_bar, 0, blockIdx.x * kRows);  // rows past the last image: zero fill
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(64, 32, true, true);
            mbar_wait(full_bar, 0);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_tile), b_addr = smem_u32(b_tile);
#pragma unroll 5
            for (int kk = 0; kk < kSteps; kk++) {
                const uint64_t da = make_smem_desc(a_addr + kk * 1024, kTile, 512, kSw128Base32);
                const uint64_t db = make_smem_desc(b_addr + kk * 1024, 4096, 512, kSw128Base32);
                umma_tf32(tmem_base, da, db, idesc, kk != 0);
            }
            umma_commit(tmem_full_bar);
        }
    } else if (warp == 4 || warp == 5) {
        // TMEM lane quarters 0 and 1 hold rows (e = 0, oc) and (e = 1, oc) in their first 16 lanes (M=64 layout)
        const int e = warp & 3;
        mbar_wait(tmem_full_bar, 0);
        tcgen05_fence_after();
        float lo[16], hi[16];
        tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * e) << 16), lo);
        tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * e) << 16) + 16u, hi);
        if (lane < 16) {
#pragma unroll
            for (int j = 0; j < 10; j++) part[(e * kOC + lane) * 10 + j] = e == 0 ? lo[j] : hi[j];
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<32>(tmem_base);
    }
    if (t < kNOut) {
        const int oc = t < kNW ? t / kTaps : t - kNW, j = t < kNW ? t % kTaps : kTaps;
        a.partials[size_t(blockIdx.x) * kNOut + t] = part[oc * 10 + j] + part[(kOC + oc) * 10 + j];
    }
    // ---- the block that finishes last adds all partials in block order (same tail as conv_bwd_weight_c1_nhwc_kernel) ----
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
        if (dst) {
            acc = a.accumulate ? *dst + acc : acc;
            *dst = acc;
        }
        if (a.sgd_w) {
            float* wdst = t < kNW ? a.sgd_w + t : a.sgd_b + (t - kNW);
            *wdst -= *a.lr_dev * acc;
        }
    }
    if (a.xc.world() > 1) {
        // compute + collective + update in one kernel: conv1's gradient never waits for a separate all-reduce launch
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

bool tc_conv1_bwd_weight_enabled() {
    static const bool on = [] {
        const char* v = getenv("B200CNN_TC_C1");  // B200CNN_TC_C1=0: CUDA-core kernel of conv_nhwc.cu (A/B timing)
        return !(v && v[0] == '0');
    }();
    return on && current_backend() != B200CNN_BACKEND_SIMT;
}

bool tc_conv1_bwd_weight_supported(const b200cnn_conv_shape* s, int max_partials) {
    return s->c == 1 && s->oc == kOC && s->k == 3 && s->stride == 2 && s->pad == 1 && s->h == 28 && s->w == 28 && s->oh == 14 &&
           s->ow == 14 && s->n >= 1 && ceil_div(s->n, kImgs) <= max_partials;
}

// x: uint8 [n][28][28]; dz: NHWC [n][14][14][16]; partials: ceil(n/2) * 160 floats; counter: one zero word.
int tc_conv1_bwd_weight(const uint8_t* x, const float* dz_nhwc, float* dw, float* db, int accumulate, float* partials,
                        unsigned* counter, const b200cnn_conv_shape* s, cudaStream_t st, float* sgd_w, float* sgd_b,
                        const float* lr_dev, const TailXchg* xchg, long w_off, long b_off) {
    if (xchg && xchg->world() > 1 && (accumulate || sgd_w || !lr_dev || !dw || !db))
        return set_error(kInvalidValue, "tc_conv1_bwd_weight: the in-kernel all-reduce needs dw/db in the gradient buffer, no accumulation, no local update");
    CUtensorMap tm;
    const uint64_t dims[2] = {32, uint64_t(s->n) * kPairs};
    const uint64_t strides[2] = {0, 128};
    const uint32_t box[2] = {32, uint32_t(kRows)};
    if (int rc = encode_tmap_f32(&tm, dz_nhwc, 2, dims, strides, box, nullptr, kTmaSw128Atom32)) return rc;
    C1Args a{x, partials, dw, db, counter, sgd_w, sgd_b, lr_dev, s->n, accumulate, xchg ? *xchg : TailXchg{}, w_off, b_off};
    static bool configured = false;
    if (!configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(tc_conv1_bwd_weight_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kSmem)));
        configured = true;
    }
    B200CNN_CUDA(launch_kernel(tc_conv1_bwd_weight_kernel, dim3(ceil_div(s->n, kImgs)), dim3(kThreads), kSmem, st, tm, a));
    B200CNN_CHECK_LAUNCH("tc_conv1_bwd_weight");
    return 0;
}

}  // namespace b200cnn
