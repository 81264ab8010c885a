// Public fully connected entry points: pick the tcgen05 path when the shape qualifies, otherwise
// the CUDA-core path.  b200cnn_set_backend() pins the choice (tests compare the two).
#include "common.cuh"
#include "backend.cuh"

using namespace b200cnn;

namespace b200cnn {
static int g_backend = B200CNN_BACKEND_AUTO;
int current_backend() { return g_backend; }
}  // namespace b200cnn

extern "C" int b200cnn_set_backend(int backend) {
    if (backend < B200CNN_BACKEND_AUTO || backend > B200CNN_BACKEND_TCGEN05) return set_error(kInvalidValue, "set_backend: unknown backend %d", backend);
    g_backend = backend;
    return 0;
}
extern "C" int b200cnn_get_backend(void) { return g_backend; }

extern "C" int b200cnn_linear_fwd(const float* x, const float* w, const float* b, float* y, int m, int n, int k,
                                  int ldy, int act, b200cnn_stream stream) {
    if (g_backend != B200CNN_BACKEND_SIMT && tc_linear_fwd_supported(m, n, k, ldy))
        return tc_linear_fwd(x, w, b, y, m, n, k, ldy, act, as_stream(stream));
    if (g_backend == B200CNN_BACKEND_TCGEN05) return set_error(kInvalidValue, "linear_fwd: shape m=%d n=%d k=%d ldy=%d not supported by the tcgen05 path", m, n, k, ldy);
    return simt_linear_fwd(x, w, b, y, m, n, k, ldy, act, as_stream(stream));
}

extern "C" int b200cnn_linear_bwd_data(const float* dz, const float* w, const float* a_prev, int act_prev, float* dx,
                                       int m, int n, int k, int lddz, b200cnn_stream stream) {
    if (g_backend != B200CNN_BACKEND_SIMT && tc_linear_bwd_data_supported(m, n, k, lddz))
        return tc_linear_bwd_data(dz, w, a_prev, act_prev, dx, m, n, k, lddz, as_stream(stream));
    if (g_backend == B200CNN_BACKEND_TCGEN05) return set_error(kInvalidValue, "linear_bwd_data: shape not supported by the tcgen05 path");
    return simt_linear_bwd_data(dz, w, a_prev, act_prev, dx, m, n, k, lddz, as_stream(stream));
}

extern "C" int b200cnn_linear_bwd_weight(const float* dz, const float* x, float* dw, float* db, int accumulate, int m,
                                         int n, int k, int lddz, b200cnn_stream stream) {
    if (g_backend != B200CNN_BACKEND_SIMT && tc_linear_bwd_weight_supported(m, n, k, lddz))
        return tc_linear_bwd_weight(dz, x, dw, db, accumulate, m, n, k, lddz, as_stream(stream));
    if (g_backend == B200CNN_BACKEND_TCGEN05) return set_error(kInvalidValue, "linear_bwd_weight: shape not supported by the tcgen05 path");
    return simt_linear_bwd_weight(dz, x, dw, db, accumulate, m, n, k, lddz, as_stream(stream));
}

extern "C" int b200cnn_linear_softmax_xent(const float* x, const float* w, const float* b, const uint8_t* labels, float* probs,
                                           float* dz, float* metrics, int m, int n, int k, int ld, b200cnn_stream stream) {
    if (g_backend != B200CNN_BACKEND_SIMT && tc_linear_softmax_supported(m, n, k, ld))
        return tc_linear_softmax_xent(x, w, b, labels, probs, dz, metrics, m, n, k, ld, as_stream(stream));
    if (g_backend == B200CNN_BACKEND_TCGEN05) return set_error(kInvalidValue, "linear_softmax_xent: shape not supported by the tcgen05 path");
    if (int rc = simt_linear_fwd(x, w, b, probs, m, n, k, ld, B200CNN_ACT_NONE, as_stream(stream))) return rc;
    return b200cnn_softmax_xent(probs, labels, probs, labels ? dz : nullptr, labels ? metrics : nullptr, m, n, ld, stream);
}

extern "C" int b200cnn_mlp_head(const float* h1, const float* w2, const float* b2, const float* w3, const float* b3,
                                const uint8_t* labels, float* h2, float* probs, float* dz3, float* dz2, float* dz1, float* metrics,
                                int m, int n1, int n2, int nc, int ldc, b200cnn_stream stream) {
    if (!tc_head_supported(m, n1, n2, nc, ldc)) return set_error(kInvalidValue, "mlp_head: need 16 <= n1, n2 <= 208 (multiples of 4), nc <= ldc <= 16");
    return tc_head(h1, w2, b2, w3, b3, labels, h2, probs, dz3, dz2, dz1, metrics, m, n1, n2, nc, ldc, as_stream(stream));
}

// ---- NHWC tensor-core convolution entry points (see include/b200cnn.h) ------------------------------
extern "C" int b200cnn_conv3x3_fwd_nhwc(const float* x_nhwc, const float* w, const float* b, float* y, int ldy,
                                        const b200cnn_conv_shape* s, int act, b200cnn_stream stream) {
    if (!s || !tc_conv3x3_fwd_supported(s)) return set_error(kInvalidValue, "conv3x3_fwd_nhwc: shape not supported by the tcgen05 path");
    return tc_conv3x3_fwd(x_nhwc, w, b, y, ldy, s, act, as_stream(stream));
}

extern "C" int b200cnn_conv3x3s2_bwd_data_nhwc(const float* dz, int dz_is_chw, const float* w, const float* a_prev_nhwc,
                                               int act_prev, float* dx_nhwc, const b200cnn_conv_shape* s,
                                               b200cnn_stream stream) {
    if (!s || !tc_conv3x3s2_bwd_data_supported(s)) return set_error(kInvalidValue, "conv3x3s2_bwd_data_nhwc: shape not supported by the tcgen05 path");
    return tc_conv3x3s2_bwd_data(dz, dz_is_chw, w, a_prev_nhwc, act_prev, dx_nhwc, s, as_stream(stream));
}

// Same forward kernel with the output left NHWC [n][oh][ow][oc] (feeds a pooling / convolution layer instead of an FC layer).
extern "C" int b200cnn_conv3x3_fwd_nhwc_to_nhwc(const float* x_nhwc, const float* w, const float* b, float* y_nhwc,
                                                const b200cnn_conv_shape* s, int act, b200cnn_stream stream) {
    if (!s || !tc_conv3x3_fwd_supported(s)) return set_error(kInvalidValue, "conv3x3_fwd_nhwc_to_nhwc: shape not supported by the tcgen05 path");
    return tc_conv3x3_fwd(x_nhwc, w, b, y_nhwc, 0, s, act, as_stream(stream), 1);
}

extern "C" int b200cnn_conv3x3s1_bwd_data_nhwc(const float* dz_nhwc, const float* w, const float* a_prev_nhwc, int act_prev,
                                               float* dx_nhwc, const b200cnn_conv_shape* s, b200cnn_stream stream) {
    if (!s || !tc_conv3x3s1_bwd_data_supported(s)) return set_error(kInvalidValue, "conv3x3s1_bwd_data_nhwc: shape not supported by the tcgen05 path");
    return tc_conv3x3s1_bwd_data(dz_nhwc, w, a_prev_nhwc, act_prev, dx_nhwc, s, as_stream(stream));
}

// dw (OIHW) / db of a convolution whose input x is NHWC (or a 1-channel image, uint8 allowed) and whose
// output gradient is NHWC (or CHW when dz_is_chw).  scratch as for b200cnn_conv2d_bwd_weight.
extern "C" int b200cnn_conv2d_bwd_weight_nhwc(const void* x, int x_is_u8, const float* dz, int dz_is_chw, float* dw, float* db,
                                              int accumulate, void* scratch, const b200cnn_conv_shape* s,
                                              b200cnn_stream stream) {
    if (!s || !scratch || !nhwc_bwd_weight_supported(s, x_is_u8)) return set_error(kInvalidValue, "conv2d_bwd_weight_nhwc: unsupported shape or missing scratch");
    return nhwc_conv_bwd_weight(x, x_is_u8, dz, dz_is_chw, dw, db, accumulate, static_cast<float*>(scratch), nullptr, s, as_stream(stream));
}
extern "C" int b200cnn_conv12_fwd(const uint8_t* x, const float* w1, const float* b1, float* a1_nhwc, const float* w2, const float* b2,
                                  float* y, int ldy, const b200cnn_conv_shape* s1, const b200cnn_conv_shape* s2, b200cnn_stream stream) {
    if (!s1 || !s2 || !x || !w1 || !b1 || !a1_nhwc || !w2 || !b2 || !y) return set_error(kInvalidValue, "conv12_fwd: null argument");
    if (!tc_conv12_fwd_supported(s1, s2) || ldy < s2->oc * s2->oh * s2->ow) return set_error(kInvalidValue, "conv12_fwd: unsupported shapes");
    return tc_conv12_fwd(x, w1, b1, a1_nhwc, B200CNN_ACT_RELU, w2, b2, y, ldy, B200CNN_ACT_RELU, s1, s2, as_stream(stream));
}
extern "C" int b200cnn_conv1_pool_fwd(const void* x, int x_is_u8, const float* w, const float* b, float* a_full, float* yp,
                                      int32_t* argmax, const b200cnn_conv_shape* s, int act, b200cnn_stream stream) {
    if (!s || !x || !w || !b || !yp || !argmax) return set_error(kInvalidValue, "conv1_pool_fwd: null argument");
    return conv1_pool_fwd(x, x_is_u8, w, b, a_full, yp, argmax, s, act, as_stream(stream));
}
extern "C" int b200cnn_conv1_pool_bwd_weight(const void* x, int x_is_u8, const float* dyp, const float* yp, const int32_t* argmax,
                                             int act, float* dw, float* db, int accumulate, void* scratch,
                                             const b200cnn_conv_shape* s, b200cnn_stream stream) {
    if (!s || !scratch || !x || !dyp || !yp || !argmax) return set_error(kInvalidValue, "conv1_pool_bwd_weight: null argument");
    return conv1_pool_bwd_weight(x, x_is_u8, dyp, yp, argmax, act, dw, db, accumulate, static_cast<float*>(scratch), nullptr, s,
                                 as_stream(stream));
}
extern "C" size_t b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(const b200cnn_conv_shape* s) {
    // + 128 bytes: completion counter of the single-channel kernel (must be zero before the first use)
    return nhwc_bwd_weight_scratch_floats(s) * sizeof(float) + 128;
}
