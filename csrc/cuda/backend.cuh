// Internal launch functions shared between the public entry points and the training engine.
#pragma once
#include <cuda_runtime.h>

#include "b200cnn.h"

namespace b200cnn {

int current_backend();

// CUDA-core path (gemm_simt.cu)
int simt_linear_fwd(const float* x, const float* w, const float* b, float* y, int m, int n, int k, int ldy, int act,
                    cudaStream_t st);
int simt_linear_bwd_data(const float* dz, const float* w, const float* a_prev, int act_prev, float* dx, int m, int n,
                         int k, int lddz, cudaStream_t st);
int simt_linear_bwd_weight(const float* dz, const float* x, float* dw, float* db, int accumulate, int m, int n, int k,
                           int lddz, cudaStream_t st);

// tcgen05 / TMEM / TMA path (tc_gemm.cu)
bool tc_linear_fwd_supported(int m, int n, int k, int ldy);
bool tc_linear_bwd_data_supported(int m, int n, int k, int lddz);
bool tc_linear_bwd_weight_supported(int m, int n, int k, int lddz);
int tc_linear_fwd(const float* x, const float* w, const float* b, float* y, int m, int n, int k, int ldy, int act,
                  cudaStream_t st);
int tc_linear_bwd_data(const float* dz, const float* w, const float* a_prev, int act_prev, float* dx, int m, int n,
                       int k, int lddz, cudaStream_t st);
// Last layer fused with the loss: probs = softmax(x W^T + b), dz = probs - onehot, metrics (n <= 16 classes).
bool tc_linear_softmax_supported(int m, int n, int k, int ld);
int tc_linear_softmax_xent(const float* x, const float* w, const float* b, const uint8_t* labels, float* probs, float* dz,
                           float* metrics, int m, int n, int k, int ld, cudaStream_t st);
// Same as tc_linear_bwd_data but column n of dx is stored at (n % perm_p) * perm_c + n / perm_p.
int tc_linear_bwd_data_permuted(const float* dz, const float* w, const float* a_prev, int act_prev, float* dx, int m, int n,
                                int k, int lddz, int perm_p, int perm_c, cudaStream_t st);
int tc_linear_bwd_weight(const float* dz, const float* x, float* dw, float* db, int accumulate, int m, int n, int k,
                         int lddz, cudaStream_t st);

// tc_head.cu: fc(tanh) -> fc(softmax) forward + loss + both backward-data GEMMs in one kernel
bool tc_head_supported(int m, int n1, int n2, int nc, int ldc);
int tc_head(const float* h1, const float* w2, const float* b2, const float* w3, const float* b3, const uint8_t* labels, float* h2,
            float* probs, float* dz3, float* dz2, float* dz1, float* metrics, int m, int n1, int n2, int nc, int ldc,
            cudaStream_t st);

// tc_chain.cu: three FC layers forward + loss + three data-gradient GEMMs in one cluster kernel
bool tc_fc_chain_supported(int m, int k1, int n1, int n2, int nc, int ldc);
bool tc_fc_chain_enabled();
int tc_fc_chain(const float* x, int act_x, const float* w1, const float* b1, const float* w2, const float* b2, const float* w3,
                const float* b3, const uint8_t* labels, float* h1, float* h2, float* probs, float* dz3, float* dz2, float* dz1,
                float* dx, float* metrics, int m, int k1, int n1, int n2, int nc, int ldc, cudaStream_t st);

// ... and only its middle (same interface as tc_head): h2, softmax / loss, dz3, dz2, dz1 from h1 in one cluster kernel
bool tc_fc_mid_supported(int m, int n1, int n2, int nc, int ldc);
bool tc_fc_mid_enabled();
int tc_fc_mid(const float* h1, const float* w2, const float* b2, const float* w3, const float* b3, const uint8_t* labels, float* h2,
              float* probs, float* dz3, float* dz2, float* dz1, float* metrics, int m, int n1, int n2, int nc, int ldc, cudaStream_t st);

// tc_conv.cu: implicit-GEMM 3x3 convolutions on NHWC activations
bool tc_conv3x3_fwd_supported(const b200cnn_conv_shape* s);
// early_w (here and below): the caller guarantees that the kernel launched just before this one on the stream does not
// write the weights (inside the engine it is always another layer's kernel), so their gather into shared memory may run
// BEFORE griddepcontrol.wait and overlap that kernel's tail.
int tc_conv3x3_fwd(const float* x_nhwc, const float* w, const float* b, float* y, int ldy, const b200cnn_conv_shape* s,
                   int act, cudaStream_t st, int nhwc_out = 0, int early_w = 0);
// stride-1 counterpart of tc_conv3x3s2_bwd_data (32 -> 16 channels, same spatial size): dz / a_prev / dx all NHWC
bool tc_conv3x3s1_bwd_data_supported(const b200cnn_conv_shape* s);
int tc_conv3x3s1_bwd_data(const float* dz_nhwc, const float* w, const float* a_prev_nhwc, int act_prev, float* dx_nhwc,
                          const b200cnn_conv_shape* s, cudaStream_t st, int early_w = 0);
// tc_conv_halo.cu: the stride-1 convolutions with ONE halo load per tile, taps by descriptor shifts (used by the two above)
bool tc_conv3x3s1_halo_fwd_supported(const b200cnn_conv_shape* s);
bool tc_conv3x3s1_halo_bwd_data_supported(const b200cnn_conv_shape* s);
int tc_conv3x3s1_halo_fwd(const float* x_nhwc, const float* w, const float* b, float* y, int ldy, const b200cnn_conv_shape* s, int act,
                          cudaStream_t st, int nhwc_out, int early_w = 0);
int tc_conv3x3s1_halo_bwd_data(const float* dz_nhwc, const float* w, const float* a_prev_nhwc, int act_prev, float* dx_nhwc,
                               const b200cnn_conv_shape* s, cudaStream_t st, int early_w = 0);
// pool.cu: NHWC max pooling for the tensor-core chain of the pooled topology
int maxpool_nhwc_fwd(const float* x, float* y, int32_t* argmax, int n, int c, int h, int w, int k, int stride, int out_chw, cudaStream_t st);
int maxpool_nhwc_bwd(const float* dy, int dy_chw, const int32_t* argmax, const float* a_prev, int act_prev, float* dx, int n, int c,
                     int h, int w, int k, int stride, cudaStream_t st);
bool tc_conv3x3s2_bwd_data_supported(const b200cnn_conv_shape* s);
int tc_conv3x3s2_bwd_data(const float* dz, int dz_is_chw, const float* w, const float* a_prev_nhwc, int act_prev,
                          float* dx_nhwc, const b200cnn_conv_shape* s, cudaStream_t st, int early_w = 0);

// tc_conv_wgrad.cu: split-K implicit-GEMM weight gradient of the 16->32 channel 3x3/s2 convolution (14x14 -> 7x7)
bool tc_conv3x3s2_bwd_weight_supported(const b200cnn_conv_shape* s);
size_t tc_conv3x3s2_bwd_weight_scratch_floats(int n);
int tc_conv3x3s2_bwd_weight(const float* x_nhwc, const float* dz, int dz_is_chw, float* dw, float* db, int accumulate,
                            float* part, const b200cnn_conv_shape* s, cudaStream_t st);

// ... and of the stride-1 16->32 channel convolution at 14x14 (pooled topology): x, dz NHWC
bool tc_conv3x3s1_bwd_weight_supported(const b200cnn_conv_shape* s);
size_t tc_conv3x3s1_bwd_weight_scratch_floats(int n);
int tc_conv3x3s1_bwd_weight(const float* x_nhwc, const float* dz_nhwc, float* dw, float* db, int accumulate, float* part,
                            const b200cnn_conv_shape* s, cudaStream_t st);

// tc_conv1_wgrad.cu: weight gradient of the 1 -> 16 channel first convolution on the tensor cores (opt-in, see the file)
bool tc_conv1_bwd_weight_enabled();
bool tc_conv1_bwd_weight_supported(const b200cnn_conv_shape* s, int max_partials);
struct TailXchg;  // csrc/comm/comm.cuh
int tc_conv1_bwd_weight(const uint8_t* x, const float* dz_nhwc, float* dw, float* db, int accumulate, float* partials,
                        unsigned* counter, const b200cnn_conv_shape* s, cudaStream_t st, float* sgd_w, float* sgd_b,
                        const float* lr_dev, const TailXchg* xchg = nullptr, long w_off = 0, long b_off = 0);

// tc_conv12.cu: forward of both convolutions of the reference topology as one kernel (the root of the captured step)
bool tc_conv12_fwd_enabled();
bool tc_conv12_fwd_supported(const b200cnn_conv_shape* s1, const b200cnn_conv_shape* s2);
int tc_conv12_fwd(const uint8_t* x, const float* w1, const float* b1, float* a1_nhwc, int act1, const float* w2, const float* b2, float* y,
                  int ldy, int act2, const b200cnn_conv_shape* s1, const b200cnn_conv_shape* s2, cudaStream_t st);

// tc_conv_tail.cu: conv2 backward-data + conv1 weight gradient (+ update / exchange) as one kernel (reference topology)
bool tc_conv_tail_enabled();
bool tc_conv_tail_supported(const b200cnn_conv_shape* s2, const b200cnn_conv_shape* s1, int max_partials);
int tc_conv2dx_conv1dw(const float* dz2, int dz_is_chw, const float* w2, const float* a_prev_nhwc, int act_prev, float* dx_nhwc,
                       const b200cnn_conv_shape* s2, const uint8_t* x, float* dw1, float* db1, float* partials, unsigned* counter,
                       const b200cnn_conv_shape* s1, float* sgd_w, float* sgd_b, const float* lr_dev, const TailXchg* xchg, long w_off,
                       long b_off, int early_w, cudaStream_t st);

// conv_nhwc.cu / conv.cu
bool nhwc_bwd_weight_supported(const b200cnn_conv_shape* s, int x_is_u8);
int nhwc_bwd_weight_chunks(int n, int* spb_out);
int nhwc_bwd_weight_chunks_for(const b200cnn_conv_shape* s, int* spb_out);
size_t nhwc_bwd_weight_smem_bytes(const b200cnn_conv_shape* s);
size_t nhwc_bwd_weight_scratch_floats(const b200cnn_conv_shape* s);
int nhwc_conv_bwd_weight(const void* x, int x_is_u8, const float* dz, int dz_chw, float* dw, float* db, int accumulate,
                         float* partials, unsigned* counter, const b200cnn_conv_shape* s, cudaStream_t st,
                         float* sgd_w = nullptr, float* sgd_b = nullptr, const float* lr_dev = nullptr,
                         const TailXchg* xchg = nullptr, long w_off = 0, long b_off = 0);
// fused first layer of the pooled topology: conv (1 channel, 3x3 / s1 / p1) + activation + 2x2 max pool, and its weight
// gradient straight from the pooled gradient
bool conv1_pool_supported(const b200cnn_conv_shape* s, int pool_k, int pool_stride);
int conv1_pool_fwd(const void* x, int x_is_u8, const float* w, const float* b, float* a_full, float* yp, int32_t* argmax,
                   const b200cnn_conv_shape* s, int act, cudaStream_t st);
int conv1_pool_bwd_weight(const void* x, int x_is_u8, const float* dyp, const float* yp, const int32_t* argmax, int act, float* dw,
                          float* db, int accumulate, float* partials, unsigned* counter, const b200cnn_conv_shape* s, cudaStream_t st,
                          float* sgd_w = nullptr, float* sgd_b = nullptr, const float* lr_dev = nullptr);
int simt_conv2d_fwd(const void* x, int x_is_u8, const float* w, const float* b, float* y, const b200cnn_conv_shape* s,
                    int act, int nhwc_out, cudaStream_t st);
int reduce_conv_partials(const float* partials, float* dw, float* db, int chunks, int nw, int nb, int accumulate,
                         cudaStream_t st);
int launch_conv2d_bwd_weight(const float* dz, const void* x, int x_is_u8, float* dw, float* db, int accumulate,
                             float* partials, const b200cnn_conv_shape* s, cudaStream_t st);

}  // namespace b200cnn
