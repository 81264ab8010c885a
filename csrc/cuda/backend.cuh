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
int tc_linear_bwd_weight(const float* dz, const float* x, float* dw, float* db, int accumulate, int m, int n, int k,
                         int lddz, cudaStream_t st);

// tc_conv.cu: implicit-GEMM 3x3 convolutions on NHWC activations
bool tc_conv3x3_fwd_supported(const b200cnn_conv_shape* s);
int tc_conv3x3_fwd(const float* x_nhwc, const float* w, const float* b, float* y, int ldy, const b200cnn_conv_shape* s,
                   int act, cudaStream_t st);
bool tc_conv3x3s2_bwd_data_supported(const b200cnn_conv_shape* s);
int tc_conv3x3s2_bwd_data(const float* dz_nhwc, const float* w, const float* a_prev_nhwc, int act_prev, float* dx_nhwc,
                          const b200cnn_conv_shape* s, cudaStream_t st);

// conv.cu
int launch_conv2d_bwd_weight(const float* dz, const void* x, int x_is_u8, float* dw, float* db, int accumulate,
                             float* partials, const b200cnn_conv_shape* s, cudaStream_t st);

}  // namespace b200cnn
