/* C ABI of the GPU library (libb200cnn_cuda.so), sm_100a only.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - tensors are fp32, activations NCHW / row-major [batch][features], conv weights
 *     [out_c][in_c][ky][kx], FC weights [out][in] (the reference's layouts:
 *     /root/reference/cnn.c:116-121, :181-197, :333);
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - functions return 0 on success or a cudaError_t value; b200cnn_last_error() describes it;
 *   - launchers only enqueue work, they never synchronise and never allocate.
 *
 * The reference has one GPU kernel (conv forward, fp64, batch 1: CUDAcnn.cu:167-195) and one host
 * wrapper that re-allocates and re-uploads per call (CUDAcnn.cu:198-218).  Everything else below
 * replaces code that the reference runs on the CPU.
 */
#ifndef B200CNN_H
#define B200CNN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200cnn_stream;

enum { B200CNN_ACT_NONE = 0, B200CNN_ACT_TANH = 1, B200CNN_ACT_RELU = 2 };

const char* b200cnn_last_error(void);
/* Compile-time facts, for tests and logs: "sm_100a", CUDA version, feature list. */
const char* b200cnn_build_info(void);

typedef struct b200cnn_conv_shape {
    int n, c, h, w;        /* input  N,C,H,W */
    int oc, oh, ow;        /* output channels and spatial size (passed in like Layer_create_conv, cnn.c:329) */
    int k, pad, stride;    /* square kernel */
} b200cnn_conv_shape;

/* Which implementation the GEMM-shaped entry points use.  AUTO = tcgen05 when the shape qualifies
 * (16-byte aligned row strides etc.), CUDA-core kernels otherwise.  Tests pin one or the other. */
enum { B200CNN_BACKEND_AUTO = 0, B200CNN_BACKEND_SIMT = 1, B200CNN_BACKEND_TCGEN05 = 2 };
int b200cnn_set_backend(int backend);
int b200cnn_get_backend(void);

/* ---- convolution (K1-K3) ------------------------------------------------------------------ */
/* y = act(conv(x, w) + b).  If x_is_u8 the input is uint8 pixels scaled by 1/255 on the fly
 * (fuses the reference's img[j]/255.0, cnn.c:457).  Replaces Layer_feedForw_conv (cnn.c:175-210)
 * and conv_forward_kernel (CUDAcnn.cu:167-195). */
int b200cnn_conv2d_fwd(const void* x, int x_is_u8, const float* w, const float* b, float* y,
                       const b200cnn_conv_shape* s, int act, b200cnn_stream stream);
/* dx = conv_transpose(dz, w) * act'(a_prev) where a_prev is the activation that produced x
 * (pass act_prev = NONE and a_prev = NULL for a raw input gradient).  Replaces the data half of
 * Layer_feedBack_conv (cnn.c:235). */
int b200cnn_conv2d_bwd_data(const float* dz, const float* w, const float* a_prev, int act_prev, float* dx,
                            const b200cnn_conv_shape* s, b200cnn_stream stream);
/* dw[oc][ic][ky][kx] (+)= sum_n,oy,ox dz * x ; db[oc] (+)= sum dz.  Two passes: per-chunk partial
 * sums into `scratch`, then a fixed-order reduction (deterministic, no atomics).
 * scratch must hold b200cnn_conv2d_bwd_weight_scratch_bytes(s) bytes.
 * Replaces the weight half of Layer_feedBack_conv (cnn.c:236, :242). */
size_t b200cnn_conv2d_bwd_weight_scratch_bytes(const b200cnn_conv_shape* s);
int b200cnn_conv2d_bwd_weight(const float* dz, const void* x, int x_is_u8, float* dw, float* db, int accumulate,
                              void* scratch, const b200cnn_conv_shape* s, b200cnn_stream stream);

/* Same result as b200cnn_conv2d_bwd_weight for callers that may not allocate device memory (a plug-in
 * inside a host application that owns all of the GPU's memory, a latency-bound loop that cannot afford
 * cudaMalloc's device synchronisation, a process with a fixed memory map).  It has no counterpart in the
 * reference.  The partial sums live in an 8 MiB array built into the library (present on every GPU that
 * loads it, until exit); the function allocates and frees nothing.  Shapes that need more scratch than
 * B200CNN_NOALLOC_SCRATCH_BYTES are refused with an error.  The first call on each device resolves the
 * array's address (do it before any stream capture); all calls on a device share the array, so calls
 * that may overlap on the GPU must be ordered by the caller. */
#define B200CNN_NOALLOC_SCRATCH_BYTES (8u << 20)
size_t b200cnn_conv2d_bwd_weight_noalloc_capacity(void);
int b200cnn_conv2d_bwd_weight_noalloc(const float* dz, const void* x, int x_is_u8, float* dw, float* db, int accumulate,
                                      const b200cnn_conv_shape* s, b200cnn_stream stream);

/* Tensor-core (tcgen05) implicit-GEMM convolutions on NHWC activations: the engine's fast path for
 * 3x3 convolutions whose input has 16 channels.  x / dz / dx / a_prev are NHWC; the forward output is
 * written per sample CHW-flattened with row stride ldy (what the next FC layer reads, cnn.c:116-121). */
int b200cnn_conv3x3_fwd_nhwc(const float* x_nhwc, const float* w, const float* b, float* y, int ldy,
                             const b200cnn_conv_shape* s, int act, b200cnn_stream stream);
/* 3x3 / stride 2 / pad 1, 32 -> 16 channels: dx = conv_transpose(dz, w) * act'(a_prev).  dz is NHWC (TMA-fed)
 * or, with dz_is_chw, the CHW gradient an FC layer produces for its flattened input (gathered in-kernel). */
int b200cnn_conv3x3s2_bwd_data_nhwc(const float* dz, int dz_is_chw, const float* w, const float* a_prev_nhwc, int act_prev,
                                    float* dx_nhwc, const b200cnn_conv_shape* s, b200cnn_stream stream);
/* The same forward kernel with an NHWC output [n][oh][ow][oc] (the pooled topology: conv -> ReLU -> max-pool), images
 * of more than 128 output pixels are cut into bands of whole rows (14x14 -> 2 x 7 rows). */
int b200cnn_conv3x3_fwd_nhwc_to_nhwc(const float* x_nhwc, const float* w, const float* b, float* y_nhwc,
                                     const b200cnn_conv_shape* s, int act, b200cnn_stream stream);
/* 3x3 / stride 1 / pad 1, 32 -> 16 channels (cnn.c:235 for the pooled topology's second convolution): dz, a_prev, dx NHWC. */
int b200cnn_conv3x3s1_bwd_data_nhwc(const float* dz_nhwc, const float* w, const float* a_prev_nhwc, int act_prev, float* dx_nhwc,
                                    const b200cnn_conv_shape* s, b200cnn_stream stream);
/* Max pooling on NHWC activations (K7): y NHWC or, with out_chw, CHW-flattened rows for a fully connected layer;
 * backward takes dy NHWC or CHW-flattened (dy_chw) and multiplies by act'(a_prev) of the layer below. */
int b200cnn_maxpool_nhwc_fwd(const float* x, float* y, int32_t* argmax, int n, int c, int h, int w, int k, int stride, int out_chw,
                             b200cnn_stream stream);
int b200cnn_maxpool_nhwc_bwd(const float* dy, int dy_chw, const int32_t* argmax, const float* a_prev, int act_prev, float* dx, int n,
                             int c, int h, int w, int k, int stride, b200cnn_stream stream);
/* dw (OIHW) / db for NHWC activations: x NHWC (or a 1-channel image, uint8 allowed), dz NHWC or CHW.
 * Register-tiled CUDA-core kernel + fixed-order reduction of per-block partial sums held in `scratch`
 * (zero the scratch once after allocating it: its first 128 bytes hold a self-resetting completion counter; the
 * buffer may then be shared by several layers as long as their calls do not overlap on the GPU). */
size_t b200cnn_conv2d_bwd_weight_nhwc_scratch_bytes(const b200cnn_conv_shape* s);
int b200cnn_conv2d_bwd_weight_nhwc(const void* x, int x_is_u8, const float* dz, int dz_is_chw, float* dw, float* db,
                                   int accumulate, void* scratch, const b200cnn_conv_shape* s, b200cnn_stream stream);
/* Forward of both convolutions of the reference topology as one kernel (conv 1->16 3x3/s2/p1 + ReLU on the CUDA cores,
 * written to a1 (NHWC [n][14][14][16]) AND straight into the tcgen05 operand tiles of conv 16->32 3x3/s2/p1 + ReLU).
 * x: uint8 [n][28][28] (scaled by 1/255); y: [n][ldy], the sample's 32x7x7 outputs CHW-flattened. */
int b200cnn_conv12_fwd(const uint8_t* x, const float* w1, const float* b1, float* a1_nhwc, const float* w2, const float* b2, float* y,
                       int ldy, const b200cnn_conv_shape* s1, const b200cnn_conv_shape* s2, b200cnn_stream stream);
/* First layer of the conv -> pool topology as one kernel: convolution (1 input channel, 3x3, stride 1, pad 1; x uint8
 * scaled by 1/255, or float) + activation + 2x2 / stride-2 max pool.  yp / argmax: NHWC [n][oh/2][ow/2][oc] (argmax = index
 * oy * ow + ox of the winning un-pooled pixel); a_full: NULL, or NHWC [n][oh][ow][oc] to also keep the un-pooled activations.
 * The weight gradient is taken straight from the POOLED gradient dyp (only window winners contribute); yp is the pooled
 * output (for act'), scratch as for b200cnn_conv2d_bwd_weight_nhwc. */
int b200cnn_conv1_pool_fwd(const void* x, int x_is_u8, const float* w, const float* b, float* a_full, float* yp, int32_t* argmax,
                           const b200cnn_conv_shape* s, int act, b200cnn_stream stream);
int b200cnn_conv1_pool_bwd_weight(const void* x, int x_is_u8, const float* dyp, const float* yp, const int32_t* argmax, int act,
                                  float* dw, float* db, int accumulate, void* scratch, const b200cnn_conv_shape* s,
                                  b200cnn_stream stream);

/* ---- fully connected (K4, K6) ------------------------------------------------------------- */
/* y[m][n] = act(sum_k x[m][k] w[n][k] + b[n]); ldy = row stride of y in floats (>= n).
 * Replaces Layer_feedForw_full (cnn.c:113-150). */
int b200cnn_linear_fwd(const float* x, const float* w, const float* b, float* y, int m, int n, int k, int ldy,
                       int act, b200cnn_stream stream);
/* dx[m][k] = (sum_n dz[m][n] w[n][k]) * act'(a_prev[m][k]); lddz = row stride of dz. (cnn.c:166) */
int b200cnn_linear_bwd_data(const float* dz, const float* w, const float* a_prev, int act_prev, float* dx, int m,
                            int n, int k, int lddz, b200cnn_stream stream);
/* dw[n][k] (+)= sum_m dz[m][n] x[m][k]; db[n] (+)= sum_m dz[m][n]. (cnn.c:167, :170) */
int b200cnn_linear_bwd_weight(const float* dz, const float* x, float* dw, float* db, int accumulate, int m, int n,
                              int k, int lddz, b200cnn_stream stream);

/* ---- softmax + cross-entropy gradient + metrics (K5) -------------------------------------- */
/* logits[m][ld] (first n columns valid) -> probs (optional, same stride), dz = probs - onehot(label)
 * (columns n..ld-1 of dz are written as 0), and metrics[4] += {sum CE loss, sum MSE of (p - y) over
 * n outputs (the reference's printed "error", cnn.c:275-282), #correct (first-max, cnn.c:508-513), m}.
 * Replaces cnn.c:125-143 and :285-286. */
int b200cnn_softmax_xent(const float* logits, const uint8_t* labels, float* probs, float* dz, float* metrics,
                         int m, int n, int ld, b200cnn_stream stream);

/* Output layer fused with the loss (K5): probs[m][ld] = softmax(x W^T + b) over the first n columns,
 * dz = probs - onehot(labels) (NULL labels: probabilities only), metrics as above.  On the tcgen05 path the
 * softmax runs in the GEMM epilogue on the accumulator row (n <= 16); otherwise two CUDA-core kernels. */
int b200cnn_linear_softmax_xent(const float* x, const float* w, const float* b, const uint8_t* labels, float* probs,
                                float* dz, float* metrics, int m, int n, int k, int ld, b200cnn_stream stream);

/* The last two layers of a step fused into one tcgen05 kernel (training only):
 *   h2 = tanh(h1 W2^T + b2); probs = softmax(h2 W3^T + b3); dz3 = probs - onehot; metrics;
 *   dz2 = (dz3 W3) * (1 - h2^2); dz1 = (dz2 W2) * (1 - h1^2).
 * Intermediate operands stay in shared memory / TMEM; h2, dz3, dz2 are also written to global memory for the
 * weight-gradient kernels.  16 <= n1, n2 <= 208 (multiples of 4), nc <= ldc <= 16. */
int b200cnn_mlp_head(const float* h1, const float* w2, const float* b2, const float* w3, const float* b3, const uint8_t* labels,
                     float* h2, float* probs, float* dz3, float* dz2, float* dz1, float* metrics, int m, int n1, int n2, int nc,
                     int ldc, b200cnn_stream stream);

/* The whole fully connected part of a training step in ONE cluster kernel (csrc/cuda/tc_chain.cu):
 *   h1 = tanh(x W1^T + b1); h2 = tanh(h1 W2^T + b2); probs = softmax(h2 W3^T + b3); dz3 = probs - onehot, metrics;
 *   dz2 = (dz3 W3) * (1 - h2^2); dz1 = (dz2 W2) * (1 - h1^2); dx = (dz1 W1) * act_x'(x).
 * x [m][k1] (k1 % 4 == 0, 32 <= k1 <= 2048), w1 [n1][k1], w2 [n2][n1], w3 [nc][n2]; 8 <= n1, n2 <= 224 (multiples of 4),
 * nc <= ldc <= 16.  Replaces Layer_feedForw_full x3 (cnn.c:113-152), the loss gradient + metric (cnn.c:275-286) and
 * the error-propagation half of Layer_feedBack_full x3 (cnn.c:158-166) for a whole mini-batch. */
int b200cnn_fc_chain(const float* x, int act_x, const float* w1, const float* b1, const float* w2, const float* b2,
                     const float* w3, const float* b3, const uint8_t* labels, float* h1, float* h2, float* probs, float* dz3,
                     float* dz2, float* dz1, float* dx, float* metrics, int m, int k1, int n1, int n2, int nc, int ldc,
                     b200cnn_stream stream);

/* The four GEMMs in the middle of that chain as one cluster kernel (same arguments as b200cnn_mlp_head): from h1 it
 * produces h2 = tanh(h1 W2^T + b2), probs / dz3 / metrics, dz2 = (dz3 W3)(1 - h2^2) and dz1 = (dz2 W2)(1 - h1^2).
 * 8 <= n1, n2 <= 208 (multiples of 4), nc <= ldc <= 16.  Replaces Layer_feedForw_full x2 (cnn.c:113-152), the loss
 * gradient + metric (cnn.c:275-286) and the error propagation of Layer_feedBack_full x2 (cnn.c:158-166). */
int b200cnn_fc_mid(const float* h1, const float* w2, const float* b2, const float* w3, const float* b3, const uint8_t* labels,
                   float* h2, float* probs, float* dz3, float* dz2, float* dz1, float* metrics, int m, int n1, int n2, int nc, int ldc,
                   b200cnn_stream stream);

/* ---- max pooling (K7; not in the reference) ------------------------------------------------ */
int b200cnn_maxpool_fwd(const float* x, float* y, int32_t* argmax, int n, int c, int h, int w, int k, int stride,
                        b200cnn_stream stream);
/* dx = route(dy) * act'(a_prev): a_prev (optional) is the pooled layer's input activation, so the
 * result is the gradient w.r.t. the PRE-activation of the layer feeding the pool. */
int b200cnn_maxpool_bwd(const float* dy, const int32_t* argmax, const float* a_prev, int act_prev, float* dx, int n,
                        int c, int h, int w, int k, int stride, b200cnn_stream stream);

/* ---- SGD (K8 with world size 1) ------------------------------------------------------------ */
/* w -= lr * g; optionally g = 0.  Replaces Layer_update (cnn.c:303-311). */
int b200cnn_sgd_update(float* w, float* g, size_t n, float lr, int zero_grad, b200cnn_stream stream);

/* ---- synthetic data on the device (K10) ---------------------------------------------------- */
/* Fills images[n][h*w] (uint8) and labels[n] with the same class-prototype + noise family as the
 * CPU generator; counter-based, so any (seed, first_index) range can be produced independently. */
int b200cnn_synthetic_batch(uint8_t* images, uint8_t* labels, int n, int h, int w, int nclass, uint64_t seed,
                            uint64_t first_index, b200cnn_stream stream);

/* Random mini-batch with replacement out of a device-resident data set (the reference draws
 * index = rand() % train_size per sample, cnn.c:455): out_images[b] = images[idx_b], same for labels;
 * idx_b is a counter-based hash of (seed, step, b). */
int b200cnn_gather_batch(const uint8_t* images, const uint8_t* labels, uint32_t n_items, int item_bytes,
                         uint8_t* out_images, uint8_t* out_labels, int batch, uint64_t seed, uint64_t step,
                         uint32_t shard_begin, uint32_t shard_len, b200cnn_stream stream);

/* ======== training engine ====================================================================
 * A network is a list of layers like the reference's linked list of Layer nodes (cnn.c:15-43):
 * input, conv (+ReLU), pool, full (+tanh; the last full layer is the softmax output).
 * The engine owns activations / gradients for up to max_batch samples and runs a whole
 * mini-batch per launch sequence; it replaces Layer_setInputs / learnOutputs / update driven
 * sample by sample from main() (cnn.c:451-474).
 */
typedef struct b200cnn_layer_desc {
    int type;                 /* B200CNN_LAYER_* */
    int depth, width, height; /* output shape */
    int k, pad, stride;       /* conv / pool */
} b200cnn_layer_desc;
enum { B200CNN_LAYER_INPUT = 0, B200CNN_LAYER_FULL = 1, B200CNN_LAYER_CONV = 2, B200CNN_LAYER_POOL = 3 };

typedef struct b200cnn_engine b200cnn_engine;

b200cnn_engine* b200cnn_engine_create(const b200cnn_layer_desc* layers, int nlayers, int max_batch);
void b200cnn_engine_destroy(b200cnn_engine* e);
/* Dense parameter count (360,810 for the reference network) and the size in floats of the padded
 * flat buffer the engine actually uses (every tensor starts on a 128-byte boundary). */
size_t b200cnn_engine_num_params(const b200cnn_engine* e);
size_t b200cnn_engine_flat_floats(const b200cnn_engine* e);
/* offsets[2*l] / [2*l+1] = float offset of layer l's weights / biases in the flat buffer (-1 if none);
 * sizes likewise. */
void b200cnn_engine_param_layout(const b200cnn_engine* e, long* offsets, long* sizes);
/* Use caller-provided device buffers (flat_floats floats each) for parameters and gradients — the
 * data-parallel path hands in NVLink-mapped symmetric memory here.  Call before the first step. */
int b200cnn_engine_bind_buffers(b200cnn_engine* e, float* params, float* grads);
float* b200cnn_engine_params(b200cnn_engine* e);
float* b200cnn_engine_grads(b200cnn_engine* e);
/* Dense host <-> device parameter transfer (layer order, weights then biases: the CPU engine's order). */
int b200cnn_engine_set_params(b200cnn_engine* e, const float* dense_host);
int b200cnn_engine_get_params(b200cnn_engine* e, float* dense_host);
int b200cnn_engine_get_grads(b200cnn_engine* e, float* dense_host);

/* Forward `batch` samples.  images: uint8 [batch][c*h*w] (x_is_u8) or fp32.  With labels the loss
 * kernel also runs (probabilities, dz of the last layer, metrics). */
int b200cnn_engine_forward(b200cnn_engine* e, const void* images, int x_is_u8, const uint8_t* labels, int batch,
                           b200cnn_stream stream);
/* Backward from the last forward: fills the gradient buffer (sum over the batch, like u_weights). */
int b200cnn_engine_backward(b200cnn_engine* e, b200cnn_stream stream);
/* params -= lr * grads (single GPU). */
int b200cnn_engine_sgd(b200cnn_engine* e, float lr, b200cnn_stream stream);
/* forward + backward + update in one call; lr is the per-sample-sum rate, i.e. rate / batch
 * (cnn.c:468).  With use_graph the launch sequence is captured once per (pointers, batch) and
 * replayed as a CUDA graph. */
int b200cnn_engine_train_step(b200cnn_engine* e, const void* images, int x_is_u8, const uint8_t* labels, int batch,
                              float lr, int use_graph, b200cnn_stream stream);
/* A step fed from pinned HOST memory: copies the inputs to one of two device staging slots on an internal copy
 * stream (so the next batch's copy overlaps this batch's training), replays the captured step on `stream` and copies
 * the 4 metrics back to `metrics_host` (pinned; may be NULL).  Never blocks the host; synchronise before reading
 * `metrics_host`.  The first call allocates the staging slots. */
int b200cnn_engine_train_step_host(b200cnn_engine* e, const void* images_host, int x_is_u8, const uint8_t* labels_host,
                                   int batch, float lr, float* metrics_host, b200cnn_stream stream);
/* forward + backward only, graph-captured like a full step; gradients stay in the gradient buffer (for update
 * paths outside the engine, e.g. the NCCL all-reduce baseline). */
int b200cnn_engine_forward_backward(b200cnn_engine* e, const void* images, int x_is_u8, const uint8_t* labels, int batch,
                                    int use_graph, b200cnn_stream stream);
/* metrics: {sum CE, sum MSE, #correct, #samples} accumulated since the last reset (device buffer). */
float* b200cnn_engine_metrics(b200cnn_engine* e);
int b200cnn_engine_reset_metrics(b200cnn_engine* e, b200cnn_stream stream);
/* Copy layer l's activations of the last forward (batch * nnodes floats; last layer: probabilities). */
int b200cnn_engine_layer_output(b200cnn_engine* e, int layer, int batch, float* host);
/* Number of kernels one train_step launches (for launch accounting in benchmarks). */
int b200cnn_engine_launches_per_step(const b200cnn_engine* e);
/* Which convolution kernels the engine runs under the current backend setting: the generic CHW kernels on the CUDA
 * cores, the NHWC chain with tcgen05 implicit GEMMs (reference topology), or the same for conv -> pool -> conv -> pool. */
#define B200CNN_CONV_PATH_GENERIC 0
#define B200CNN_CONV_PATH_TC 1
#define B200CNN_CONV_PATH_TC_POOL 2
int b200cnn_engine_conv_path(const b200cnn_engine* e);

#ifdef __cplusplus
}
#endif
#endif
