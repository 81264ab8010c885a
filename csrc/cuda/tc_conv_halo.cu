// 3x3 / stride 1 / pad 1 convolutions on the tensor cores with ONE operand load per tile.
//
// tc_conv.cu builds the implicit GEMM from nine TMA boxes, one per filter tap: the same pixels cross L2 -> shared
// memory nine times (58 MB per step for the pooled topology's 14x14x32 gradient) and a CTA needs 9 x 16 KiB of
// shared memory, so one CTA per SM and nothing else beside it.  For stride 1 the nine operands are the SAME pixels
// shifted by whole pixels, and in NHWC a pixel is exactly one swizzle row (32 floats = 128 B, or 16 floats = 64 B).
// So here a CTA loads its band of the image ONCE, with a one-pixel halo and padded to a power-of-two width
// (TMA zero-fills everything outside the image):
//     tile row  (yl + 1) * PW + (x + 1)   <->   pixel (y0 + yl, x),   PW = 16 for widths <= 14 (32 for <= 30)
// and runs the GEMM over M = "padded" output pixels m = yl * PW + x.  A K-major swizzled tile is linear in its row
// index (row r lives at r * pitch; the 16-byte chunks inside a row are XOR-ed with address bits), so the operand of
// tap (ky, kx) is the same tile with the descriptor's start address moved by (ky * PW + kx) rows — no data moves.
// Measured on B200: tcgen05.mma applies the swizzle XOR to the ABSOLUTE shared-memory address bits, exactly as TMA
// does when it writes the tile, so a start address that is not aligned to the 8-row swizzle period needs nothing
// else — the descriptor's base-offset field must stay 0 (with it set to the period phase, (addr >> 7) & 7, every
// result is wrong; tests/test_gpu_ops.py::test_tcgen05_conv3x3_stride1_*).  The PW - W columns at the right of every row compute garbage that the epilogue drops
// (12.5 % of the MMA rows at width 14): 20 KiB (gradient) / 10 KiB (activations) of operand per CTA instead of
// 144 / 72 KiB, five CTAs per SM, one TMA instruction instead of nine.
//
//   forward        y = act(b + conv(x, w))          x NHWC [n][h][w][16]  (64-byte rows, SW64)
//   backward-data  dx = act'(a_prev) * conv(dz, mirrored w)      dz NHWC [n][h][w][32]  (128-byte rows, SW128)
//
// Replaces Layer_feedForw_conv and the data half of Layer_feedBack_conv (/root/reference/cnn.c:175-247) for the
// second convolution of the conv -> pool -> conv -> pool topology.
#include <cstdlib>

#include "backend.cuh"
#include "common.cuh"
#include "tc_common.cuh"

namespace b200cnn {

namespace {

using namespace tc;

constexpr int kThreads = 192;  // warp 0: TMA, warp 1: TMEM + MMA, warps 2-5: weights gather + epilogue
constexpr int kTaps = 9;
constexpr int kRows = 128;     // UMMA M = padded output pixels per CTA

struct HaloArgs {
    const float* w;       // OIHW [oc][c][3][3]
    const float* bias;    // forward: [oc]
    const float* a_prev;  // backward: NHWC [n][h][w][NO] activations below (for act'), or NULL
    float* out;           // forward: NHWC [n][h][w][NO] or CHW-flattened rows [n][ldy]; backward: NHWC [n][h][w][NO]
    int ldy, nhwc_out;
    int n, h, w_in;
    int pw, br, bands;    // padded width, image rows per band (pw * br = 128), bands per image
    int act;
    int early_w;          // gather the weights before griddepcontrol.wait (see backend.cuh)
};

// KC: channels per pixel of the loaded operand (GEMM K per tap), NO: channels produced (GEMM N).
template <int KC, int NO, bool kBwd>
__global__ void __launch_bounds__(kThreads)
tc_conv3x3s1_halo_kernel(const __grid_constant__ CUtensorMap tmap_in, HaloArgs a) {
    static_assert(KC == 16 || KC == 32, "one 64- or 128-byte swizzle row per pixel");
    constexpr int pitch = KC * 4;
    constexpr int kBTile = NO * pitch;  // per tap: NO rows (output channels) x KC floats
    constexpr int kTmemCols = NO < 32 ? 32 : NO;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int tile_rows = (a.br + 2) * a.pw;
    const int a_bytes = (tile_rows * pitch + 1023) & ~1023;
    uint8_t* a_tile = smem;
    uint8_t* b_tiles = smem + a_bytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(b_tiles + kTaps * kBTile);
    uint64_t* tmem_full_bar = full_bar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    const int band = int(blockIdx.x) % a.bands, n0 = int(blockIdx.x) / a.bands;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_in);
        mbar_init(full_bar, 1);
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Weights -> per-tap K-major tiles in the operand's swizzle; every load of a thread in flight at once.
    auto gather_weights = [&]() {
        constexpr int kW = NO * KC * kTaps, kStride = kThreads - 64, kBatch = 12;
        for (int i0 = threadIdx.x - 64; i0 < kW; i0 += kBatch * kStride) {  // i walks OIHW linearly
            float v[kBatch];
#pragma unroll
            for (int u = 0; u < kBatch; u++) v[u] = (i0 + u * kStride < kW) ? __ldg(a.w + i0 + u * kStride) : 0.f;
#pragma unroll
            for (int u = 0; u < kBatch; u++) {
                const int i = i0 + u * kStride;
                if (i >= kW) break;
                const int t = i % kTaps;
                // forward: OIHW = [NO][KC][9], tile row = oc, K = c.   backward: OIHW = [KC][NO][9], tile row = c, K = oc.
                const int row = kBwd ? (i / kTaps) % NO : i / (kTaps * KC);
                const int k = kBwd ? i / (kTaps * NO) : (i / kTaps) % KC;
                const int chunk = KC == 32 ? ((k >> 2) ^ (row & 7)) : ((k >> 2) ^ ((row >> 1) & 3));
                *reinterpret_cast<float*>(b_tiles + t * kBTile + row * pitch + chunk * 16 + (k & 3) * 4) = v[u];
            }
        }
        fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor-core (async) proxy
    };
    if (a.early_w && warp >= 2) gather_weights();  // overlaps the tail of the kernel in front
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0) {
            mbar_arrive_expect_tx(full_bar, uint32_t(tile_rows * pitch));
            tma_load_4d(a_tile, &tmap_in, full_bar, 0, -1, band * a.br - 1, n0);  // the band with its halo, zero outside the image
        }
    } else if (warp >= 2 && !a.early_w) {
        gather_weights();
    }
    __syncthreads();

    if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(kRows, NO, false, false);
            mbar_wait(full_bar, 0);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_tile), b_addr = smem_u32(b_tiles);
#pragma unroll
            for (int t = 0; t < kTaps; t++) {
                const int ky = t / 3, kx = t % 3;
                // forward: pixel (y - 1 + ky, x - 1 + kx); backward (mirrored taps): pixel (y + 1 - ky, x + 1 - kx)
                const int shift = kBwd ? (2 - ky) * a.pw + (2 - kx) : ky * a.pw + kx;
#pragma unroll
                for (int kk = 0; kk < KC / 8; kk++)
                    umma_tf32(tmem_base, make_smem_desc(a_addr + shift * pitch + kk * 32, 16, 8 * pitch, KC == 32 ? kSw128 : kSw64),
                              make_smem_desc(b_addr + t * kBTile + kk * 32, 16, 8 * pitch, KC == 32 ? kSw128 : kSw64), idesc,
                              (t | kk) != 0);
            }
            umma_commit(tmem_full_bar);
        }
    } else if (warp >= 2) {
        const int q = warp & 3;
        const int r = 32 * q + lane;
        const int yl = r / a.pw, x = r % a.pw, y = band * a.br + yl;
        const bool live = x < a.w_in && y < a.h;
        const size_t pixel = (size_t(n0) * a.h + y) * a.w_in + x;
        float pre[NO];  // forward: the bias; backward: the activations below (for act'), fetched before the accumulator is ready
        if (!kBwd) {
#pragma unroll
            for (int i = 0; i < NO; i++) pre[i] = __ldg(a.bias + i);
        } else if (live && a.a_prev) {
#pragma unroll
            for (int k4 = 0; k4 < NO / 4; k4++) {
                const float4 m4 = __ldg(reinterpret_cast<const float4*>(a.a_prev + pixel * NO) + k4);
                pre[4 * k4] = m4.x, pre[4 * k4 + 1] = m4.y, pre[4 * k4 + 2] = m4.z, pre[4 * k4 + 3] = m4.w;
            }
        }
        mbar_wait(tmem_full_bar, 0);
        tcgen05_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < NO; c0 += 16) {
            float v[16];
            __syncwarp();
            tmem_ld_32x32b_x16(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(c0), v);
            if (!live) continue;
            if (!kBwd) {
#pragma unroll
                for (int i = 0; i < 16; i++) v[i] = apply_act(a.act, v[i] + pre[c0 + i]);
            } else if (a.a_prev) {
#pragma unroll
                for (int i = 0; i < 16; i++) v[i] *= act_grad_from_output(a.act, pre[c0 + i]);
            }
            if (kBwd || a.nhwc_out) {
                float4* dp = reinterpret_cast<float4*>(a.out + pixel * NO + c0);
#pragma unroll
                for (int k4 = 0; k4 < 4; k4++) dp[k4] = make_float4(v[4 * k4], v[4 * k4 + 1], v[4 * k4 + 2], v[4 * k4 + 3]);
            } else {
                const int pix = a.h * a.w_in;
                float* op = a.out + size_t(n0) * a.ldy + size_t(c0) * pix + y * a.w_in + x;
#pragma unroll
                for (int i = 0; i < 16; i++) op[size_t(i) * pix] = v[i];
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

int padded_width(int w) { return w <= 14 ? 16 : w <= 30 ? 32 : 0; }

int halo_mode() {
    // B200CNN_CONV_HALO=0: the nine-box kernels of tc_conv.cu (A/B timing)
    static const int mode = [] {
        const char* v = getenv("B200CNN_CONV_HALO");
        return v ? atoi(v) : 1;
    }();
    return mode;
}

template <int KC, int NO, bool kBwd>
int launch_halo(const float* in_nhwc, const HaloArgs& a0, cudaStream_t st) {
    HaloArgs a = a0;
    a.pw = padded_width(a.w_in);
    a.br = kRows / a.pw;
    a.bands = ceil_div(a.h, a.br);
    CUtensorMap tm;
    const uint64_t dims[4] = {uint64_t(KC), uint64_t(a.w_in), uint64_t(a.h), uint64_t(a.n)};
    const uint64_t strides[4] = {0, uint64_t(KC) * 4, uint64_t(a.w_in) * KC * 4, uint64_t(a.h) * a.w_in * KC * 4};
    const uint32_t box[4] = {uint32_t(KC), uint32_t(a.pw), uint32_t(a.br + 2), 1};
    if (int rc = encode_tmap_f32(&tm, in_nhwc, 4, dims, strides, box, nullptr, KC == 32 ? 128 : 64)) return rc;
    auto kern = tc_conv3x3s1_halo_kernel<KC, NO, kBwd>;
    const size_t smem = size_t(((a.br + 2) * a.pw * KC * 4 + 1023) & ~1023) + size_t(kTaps) * NO * KC * 4 + 1024 + 64;
    static size_t configured = 0;
    if (smem > configured) {
        B200CNN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        configured = smem;
    }
    B200CNN_CUDA(launch_kernel(kern, dim3(a.n * a.bands), dim3(kThreads), smem, st, tm, a));
    B200CNN_CHECK_LAUNCH("tc_conv3x3s1_halo");
    return 0;
}

}  // namespace

bool tc_conv3x3s1_halo_fwd_supported(const b200cnn_conv_shape* s) {
    return halo_mode() != 0 && s->k == 3 && s->stride == 1 && s->pad == 1 && s->c == 16 && (s->oc == 16 || s->oc == 32) && s->oh == s->h &&
           s->ow == s->w && padded_width(s->w) > 0;
}
bool tc_conv3x3s1_halo_bwd_data_supported(const b200cnn_conv_shape* s) {
    return halo_mode() != 0 && s->k == 3 && s->stride == 1 && s->pad == 1 && s->c == 16 && s->oc == 32 && s->oh == s->h && s->ow == s->w &&
           padded_width(s->w) > 0;
}

// x: NHWC [n][h][w][16]; y: NHWC [n][h][w][oc] (nhwc_out) or [n][ldy] with the sample's outputs CHW-flattened.
int tc_conv3x3s1_halo_fwd(const float* x_nhwc, const float* w, const float* b, float* y, int ldy, const b200cnn_conv_shape* s, int act,
                          cudaStream_t st, int nhwc_out, int early_w) {
    if (!tc_conv3x3s1_halo_fwd_supported(s)) return set_error(kInvalidValue, "tc_conv3x3s1_halo_fwd: unsupported shape");
    HaloArgs a{};
    a.w = w, a.bias = b, a.out = y, a.ldy = ldy, a.nhwc_out = nhwc_out, a.n = s->n, a.h = s->h, a.w_in = s->w, a.act = act, a.early_w = early_w;
    return s->oc == 16 ? launch_halo<16, 16, false>(x_nhwc, a, st) : launch_halo<16, 32, false>(x_nhwc, a, st);
}

// dz: NHWC [n][h][w][32]; a_prev / dx: NHWC [n][h][w][16].
int tc_conv3x3s1_halo_bwd_data(const float* dz_nhwc, const float* w, const float* a_prev_nhwc, int act_prev, float* dx_nhwc,
                               const b200cnn_conv_shape* s, cudaStream_t st, int early_w) {
    if (!tc_conv3x3s1_halo_bwd_data_supported(s)) return set_error(kInvalidValue, "tc_conv3x3s1_halo_bwd_data: unsupported shape");
    HaloArgs a{};
    a.w = w, a.a_prev = a_prev_nhwc, a.out = dx_nhwc, a.n = s->n, a.h = s->h, a.w_in = s->w, a.act = act_prev, a.early_w = early_w;
    return launch_halo<32, 16, true>(dz_nhwc, a, st);
}

}  // namespace b200cnn
