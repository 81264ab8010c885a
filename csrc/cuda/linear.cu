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

// ---- NHWC tensor-core convolution entry points (see include/b200cnn.h) ------------------------------
extern "C" int b200cnn_conv3x3_fwd_nhwc(const float* x_nhwc, const float* w, const float* b, float* y, int ldy,
                                        const b200cnn_conv_shape* s, int act, b200cnn_stream stream) {
    if (!s || !tc_conv3x3_fwd_supported(s)) return set_error(kInvalidValue, "conv3x3_fwd_nhwc: shape not supported by the tcgen05 path");
    return tc_conv3x3_fwd(x_nhwc, w, b, y, ldy, s, act, as_stream(stream));
}

extern "C" int b200cnn_conv3x3s2_bwd_data_nhwc(const float* dz_nhwc, const float* w, const float* a_prev_nhwc, int act_prev,
                                               float* dx_nhwc, const b200cnn_conv_shape* s, b200cnn_stream stream) {
    if (!s || !tc_conv3x3s2_bwd_data_supported(s)) return set_error(kInvalidValue, "conv3x3s2_bwd_data_nhwc: shape not supported by the tcgen05 path");
    return tc_conv3x3s2_bwd_data(dz_nhwc, w, a_prev_nhwc, act_prev, dx_nhwc, s, as_stream(stream));
}
