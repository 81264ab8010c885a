// Softmax + cross-entropy gradient + the reference's metrics, one pass over the logits.
//
// Reference being replaced (all CPU): the softmax of the last layer (/root/reference/cnn.c:125-143),
// errors = outputs - onehot (cnn.c:285-286), the printed mean-squared "error" (cnn.c:275-282) and the
// first-max arg-max of the test loop (cnn.c:508-513).  The softmax max starts at -inf here (the
// reference starts it at -1).
#include <cfloat>

#include "common.cuh"

namespace b200cnn {
namespace {

constexpr int kLossThreads = 128;
constexpr int kMaxClasses = 64;

// One thread per sample (n <= 64 classes: the whole row lives in registers/local memory).
__global__ void __launch_bounds__(kLossThreads)
softmax_xent_kernel(const float* __restrict__ logits, const uint8_t* __restrict__ labels, float* __restrict__ probs,
                    float* __restrict__ dz, float* __restrict__ metrics, int m, int n, int ld) {
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    float ce = 0.0f, mse = 0.0f, correct = 0.0f, count = 0.0f;
    if (row < m) {
        const float* z = logits + size_t(row) * ld;
        float v[kMaxClasses];
        float mx = -FLT_MAX;
        int arg = 0;
        for (int j = 0; j < n; j++) {
            v[j] = z[j];
            if (mx < v[j]) {  // strict '<' keeps the first maximum, like cnn.c:510
                mx = v[j];
                arg = j;
            }
        }
        float sum = 0.0f;
        for (int j = 0; j < n; j++) {
            v[j] = __expf(v[j] - mx);
            sum += v[j];
        }
        const float inv = 1.0f / sum;
        const int label = labels ? int(labels[row]) : -1;
        for (int j = 0; j < n; j++) {
            const float p = v[j] * inv;
            const float e = p - (j == label ? 1.0f : 0.0f);
            if (probs) probs[size_t(row) * ld + j] = p;
            if (dz) dz[size_t(row) * ld + j] = e;
            mse += e * e;
            if (j == label) ce = -__logf(fmaxf(p, 1e-30f));
        }
        for (int j = n; j < ld; j++) {
            if (probs) probs[size_t(row) * ld + j] = 0.0f;
            if (dz) dz[size_t(row) * ld + j] = 0.0f;
        }
        mse /= n;
        correct = (arg == label) ? 1.0f : 0.0f;
        count = 1.0f;
    }
    if (!metrics) return;
    ce = warp_sum(ce);
    mse = warp_sum(mse);
    correct = warp_sum(correct);
    count = warp_sum(count);
    __shared__ float red[4][kLossThreads / 32];
    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    if (lane == 0) {
        red[0][warp] = ce;
        red[1][warp] = mse;
        red[2][warp] = correct;
        red[3][warp] = count;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        float s = 0.0f;
        for (int i = 0; i < kLossThreads / 32; i++) s += red[threadIdx.x][i];
        atomicAdd(metrics + threadIdx.x, s);
    }
}

}  // namespace
}  // namespace b200cnn

using namespace b200cnn;

extern "C" int b200cnn_softmax_xent(const float* logits, const uint8_t* labels, float* probs, float* dz,
                                    float* metrics, int m, int n, int ld, b200cnn_stream stream) {
    if (m <= 0 || n <= 0 || n > kMaxClasses || ld < n)
        return set_error(kInvalidValue, "softmax_xent: need 0 < n <= %d and ld >= n (got m=%d n=%d ld=%d)", kMaxClasses, m, n, ld);
    softmax_xent_kernel<<<ceil_div(m, kLossThreads), kLossThreads, 0, as_stream(stream)>>>(logits, labels, probs, dz,
                                                                                          metrics, m, n, ld);
    B200CNN_CHECK_LAUNCH("softmax_xent");
    return 0;
}
