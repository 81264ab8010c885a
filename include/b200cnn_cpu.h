/* C ABI of the CPU engine (libb200cnn_cpu.so).
 *
 * Everything here is fp64 and single-sample, like the reference's Layer API
 * (/root/reference/cnn.c:249-314).  It exists so that other languages (the Python package uses
 * ctypes) can drive the native engine: as a numerical oracle for the GPU kernels and for the
 * serial / data-parallel CPU trainers.
 */
#ifndef B200CNN_CPU_H
#define B200CNN_CPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200cnn_cpu_net b200cnn_cpu_net;

/* Layer kinds, numbered like the reference's LayerType (cnn.c:8-12) plus a pooling kind. */
enum { B200CNN_LAYER_INPUT = 0, B200CNN_LAYER_FULL = 1, B200CNN_LAYER_CONV = 2, B200CNN_LAYER_POOL = 3 };

/* quirks: bit0 = shared conv slice (D1), bit1 = early update (D2), bit2 = softmax max starts at -1 (D4). */
b200cnn_cpu_net* b200cnn_cpu_net_create(unsigned quirks);
void b200cnn_cpu_net_destroy(b200cnn_cpu_net* net);

/* Constructors, cf. Layer_create_input/_conv/_full (cnn.c:316-342).  Weights are drawn from the
 * net's glibc-compatible generator, seeded by b200cnn_cpu_net_seed (srand(0) equivalent: seed 0). */
void b200cnn_cpu_net_seed(b200cnn_cpu_net* net, unsigned seed);
int b200cnn_cpu_net_add_input(b200cnn_cpu_net* net, int depth, int width, int height);
int b200cnn_cpu_net_add_conv(b200cnn_cpu_net* net, int depth, int width, int height, int kernsize, int padding,
                             int stride, double std);
int b200cnn_cpu_net_add_full(b200cnn_cpu_net* net, int nnodes, double std);
int b200cnn_cpu_net_add_pool(b200cnn_cpu_net* net, int kernsize, int stride);

int b200cnn_cpu_net_num_layers(const b200cnn_cpu_net* net);
/* info[8] = {type, depth, width, height, kernsize, padding, stride, nweights}; returns nbiases. */
int b200cnn_cpu_net_layer_info(const b200cnn_cpu_net* net, int layer, int* info);
size_t b200cnn_cpu_net_num_params(const b200cnn_cpu_net* net);

/* Layer_setInputs / getOutputs / learnOutputs / getErrorTotal / update (cnn.c:249-314). */
void b200cnn_cpu_net_set_inputs(b200cnn_cpu_net* net, const double* x);
void b200cnn_cpu_net_get_outputs(const b200cnn_cpu_net* net, double* y);
void b200cnn_cpu_net_learn_outputs(b200cnn_cpu_net* net, const double* target);
double b200cnn_cpu_net_error_total(const b200cnn_cpu_net* net);
/* Mini-batch schedule (no counterpart in the reference, which only has the per-sample calls above): the same
 * accumulation into the update buffers as n x (set_inputs + learn_outputs), with the fully connected layers run
 * as matrix-matrix loops and the convolutions as row gathers; sample_error[n] (may be NULL) receives what
 * error_total() would have returned after each sample.  Falls back to the per-sample calls in quirks mode. */
int b200cnn_cpu_net_batch_supported(const b200cnn_cpu_net* net);
void b200cnn_cpu_net_learn_batch(b200cnn_cpu_net* net, const double* xs, const double* targets, int n, double* sample_error);
void b200cnn_cpu_net_forward_batch(b200cnn_cpu_net* net, const double* xs, int n, double* ys);
void b200cnn_cpu_net_update(b200cnn_cpu_net* net, double rate);

/* Flat views, layer order, weights then biases per layer. */
void b200cnn_cpu_net_get_params(const b200cnn_cpu_net* net, double* out);
void b200cnn_cpu_net_set_params(b200cnn_cpu_net* net, const double* in);
void b200cnn_cpu_net_get_updates(const b200cnn_cpu_net* net, double* out);
/* which: 0 outputs, 1 gradients, 2 errors.  Copies nnodes doubles; returns nnodes. */
int b200cnn_cpu_net_layer_array(const b200cnn_cpu_net* net, int layer, int which, double* out);

int b200cnn_cpu_net_save(const b200cnn_cpu_net* net, const char* path);
int b200cnn_cpu_net_load(b200cnn_cpu_net* net, const char* path);

/* next `n` values of glibc rand() after srand(seed). */
void b200cnn_glibc_rand(unsigned seed, int n, int* out);
/* next rand() value of the net's own stream (the trainer samples from it after init, cnn.c:455). */
int b200cnn_cpu_net_rand(b200cnn_cpu_net* net);

/* Synthetic MNIST-shaped data: images n*h*w uint8, labels n uint8. */
void b200cnn_make_synthetic(unsigned n, unsigned h, unsigned w, unsigned nclass, uint64_t seed, uint8_t* images,
                            uint8_t* labels);
/* IDX files: returns 0 on success, 111 on failure (the reference's file-error code). */
int b200cnn_idx_write(const char* path, int ndims, const uint32_t* dims, const uint8_t* data);
/* First call with data == NULL to learn ndims/dims (dims has room for 4), then again with a buffer. */
int b200cnn_idx_read(const char* path, int* ndims, uint32_t* dims, uint8_t* data, size_t capacity);

#ifdef __cplusplus
}
#endif
#endif
