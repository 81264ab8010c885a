// Generic-shape fp32 GEMM on CUDA cores with the fully connected layers' fused epilogues.  This is
// the any-shape path of the library; tc_gemm.cu (tcgen05 / TMEM / TMA) takes over for shapes whose
// strides satisfy TMA's 16-byte rules.
//
// Reference being replaced: Layer_feedForw_full / Layer_feedBack_full (/root/reference/cnn.c:113-173),
// which run on the CPU in all four reference programs.
#include "common.cuh"

namespace b200cnn {

namespace {

constexpr int BM = 64, BN = 64, BK = 16, TM = 4, TN = 4;
constexpr int kGemmThreads = (BM / TM) * (BN / TN);  // 256

enum Epilogue { kEpiBiasAct = 0, kEpiActGrad = 1, kEpiStore = 2 };

// C[m][n] = sum_k A(m,k) * B(k,n) with A(m,k) = A[m*a_m + k*a_k], B(k,n) = B[k*b_k + n*b_n].
struct GemmArgs {
    const float* A;
    const float* B;
    float* C;
    int M, N, K;
    long a_m, a_k, b_k, b_n;
    int ldc;
    // epilogue operands
    const float* bias;  // kEpiBiasAct: per-column bias
    const float* aux;   // kEpiActGrad: activation output with C's shape (row stride ld_aux)
    int ld_aux;
    int act;
    float* rowsum;      // kEpiStore: rowsum[m] (+)= sum_k A(m,k)   (bias gradient), or NULL
    int accumulate;     // kEpiStore: C += instead of C =
};

template <int EPI>
__global__ void __launch_bounds__(kGemmThreads) gemm_simt_kernel(GemmArgs g) {
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN + 4];
    const int tid = threadIdx.x;
    const int tx = tid % (BN / TN), ty = tid / (BN / TN);
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; i++)
#pragma unroll
        for (int j = 0; j < TN; j++) acc[i][j] = 0.0f;
    float rs = 0.0f;  // row-sum of A for row (m0 + tid) when EPI == kEpiStore and this is column-block 0

    // Thread -> element mapping for the tile loads follows whichever dimension is contiguous.
    const bool a_kfast = (g.a_k == 1), b_kfast = (g.b_k == 1);

    for (int k0 = 0; k0 < g.K; k0 += BK) {
#pragma unroll
        for (int i = 0; i < (BM * BK) / kGemmThreads; i++) {
            const int e = tid + i * kGemmThreads;
            const int kk = a_kfast ? e % BK : e / BM, mm = a_kfast ? e / BK : e % BM;
            const int gm = m0 + mm, gk = k0 + kk;
            As[kk][mm] = (gm < g.M && gk < g.K) ? __ldg(g.A + gm * g.a_m + gk * g.a_k) : 0.0f;
        }
#pragma unroll
        for (int i = 0; i < (BN * BK) / kGemmThreads; i++) {
            const int e = tid + i * kGemmThreads;
            const int kk = b_kfast ? e % BK : e / BN, nn = b_kfast ? e / BK : e % BN;
            const int gn = n0 + nn, gk = k0 + kk;
            Bs[kk][nn] = (gn < g.N && gk < g.K) ? __ldg(g.B + gk * g.b_k + gn * g.b_n) : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; kk++) {
            float a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM; i++) a[i] = As[kk][ty * TM + i];
#pragma unroll
            for (int j = 0; j < TN; j++) b[j] = Bs[kk][tx * TN + j];
#pragma unroll
            for (int i = 0; i < TM; i++)
#pragma unroll
                for (int j = 0; j < TN; j++) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (EPI == kEpiStore && g.rowsum && blockIdx.x == 0 && tid < BM) {
#pragma unroll
            for (int kk = 0; kk < BK; kk++) rs += As[kk][tid];
        }
        __syncthreads();
    }

#pragma unroll
    for (int i = 0; i < TM; i++) {
        const int m = m0 + ty * TM + i;
        if (m >= g.M) continue;
#pragma unroll
        for (int j = 0; j < TN; j++) {
            const int n = n0 + tx * TN + j;
            if (n >= g.N) continue;
            float v = acc[i][j];
            if (EPI == kEpiBiasAct) {
                if (g.bias) v += __ldg(g.bias + n);
                v = apply_act(g.act, v);
            } else if (EPI == kEpiActGrad) {
                if (g.aux) v *= act_grad_from_output(g.act, __ldg(g.aux + size_t(m) * g.ld_aux + n));
            } else if (g.accumulate) {
                v += g.C[size_t(m) * g.ldc + n];
            }
            g.C[size_t(m) * g.ldc + n] = v;
        }
    }
    if (EPI == kEpiStore && g.rowsum && blockIdx.x == 0 && tid < BM && m0 + tid < g.M)
        g.rowsum[m0 + tid] = g.accumulate ? g.rowsum[m0 + tid] + rs : rs;
}

template <int EPI>
int launch(const GemmArgs& g, cudaStream_t st, const char* name) {
    if (g.M <= 0 || g.N <= 0 || g.K <= 0) return set_error(kInvalidValue, "%s: empty problem", name);
    const dim3 grid(ceil_div(g.N, BN), ceil_div(g.M, BM));
    gemm_simt_kernel<EPI><<<grid, kGemmThreads, 0, st>>>(g);
    B200CNN_CHECK_LAUNCH(name);
    return 0;
}

}  // namespace

int simt_linear_fwd(const float* x, const float* w, const float* b, float* y, int m, int n, int k, int ldy, int act,
                    cudaStream_t st) {
    GemmArgs g{};
    g.A = x; g.a_m = k; g.a_k = 1;
    g.B = w; g.b_k = 1; g.b_n = k;
    g.C = y; g.ldc = ldy;
    g.M = m; g.N = n; g.K = k;
    g.bias = b; g.act = act;
    return launch<kEpiBiasAct>(g, st, "linear_fwd");
}

int simt_linear_bwd_data(const float* dz, const float* w, const float* a_prev, int act_prev, float* dx, int m, int n,
                         int k, int lddz, cudaStream_t st) {
    GemmArgs g{};
    g.A = dz; g.a_m = lddz; g.a_k = 1;   // [m][n], reduce over n
    g.B = w; g.b_k = k; g.b_n = 1;       // w[n][k]: B(kk=n, nn=k)
    g.C = dx; g.ldc = k;
    g.M = m; g.N = k; g.K = n;
    g.aux = a_prev; g.ld_aux = k; g.act = act_prev;
    return launch<kEpiActGrad>(g, st, "linear_bwd_data");
}

int simt_linear_bwd_weight(const float* dz, const float* x, float* dw, float* db, int accumulate, int m, int n, int k,
                           int lddz, cudaStream_t st) {
    GemmArgs g{};
    g.A = dz; g.a_m = 1; g.a_k = lddz;   // A(row=n, red=m) = dz[m][n]
    g.B = x; g.b_k = k; g.b_n = 1;       // B(red=m, col=k) = x[m][k]
    g.C = dw; g.ldc = k;
    g.M = n; g.N = k; g.K = m;
    g.rowsum = db; g.accumulate = accumulate;
    return launch<kEpiStore>(g, st, "linear_bwd_weight");
}

}  // namespace b200cnn
