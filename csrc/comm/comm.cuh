// Internal interface of the NVLink communicator (see include/b200cnn_comm.h for the C ABI).
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

#include "b200cnn_comm.h"
#include "ll.cuh"

struct b200cnn_comm;

namespace b200cnn {

// What a kernel OUTSIDE the communicator needs to exchange a small slice with its peers itself (the weight-gradient
// kernel of the first convolution does its own one-shot all-reduce + SGD in the block that finishes last:
// csrc/cuda/tc_conv1_wgrad.cu, csrc/comm/xchg_device.cuh).  Filled by comm_tail_xchg() for the step being enqueued.
struct TailXchg {
    ll::Channel chan;       // push-protocol channel 2 (ll.cuh): world, rank, peers' regions, counters, error flag
    size_t params_off = 0;  // byte offset of the parameters inside a region
    __host__ __device__ int world() const { return chan.world; }
};
void comm_tail_xchg(const b200cnn_comm* c, TailXchg* out);

// Enqueue the fused all-reduce + 1/world + SGD for floats [offset, offset + count) of the flat
// parameter / gradient buffers.  The kernel runs on the communicator's own stream, ordered after
// everything enqueued on `after` so far (event fork), so it overlaps whatever `after` does next.
int comm_allreduce_sgd_slice(b200cnn_comm* c, size_t offset, size_t count, const float* lr_dev, cudaStream_t after);
// The same update enqueued on `st` itself, on a second flag channel (may overlap a slice on the communicator's stream).
int comm_allreduce_sgd_inline(b200cnn_comm* c, size_t offset, size_t count, const float* lr_dev, cudaStream_t st);
// Deferred second half of the two-shot push calls (see allreduce_sgd.cu): enqueue the wait at the start of a step,
// consume it before the first kernel that reads updated parameters.
int comm_begin_step(b200cnn_comm* c, cudaStream_t st);
// Would an update of `count` floats on push channel `ch` leave peers' stores pending when its kernel returns?
bool comm_update_is_deferred(const b200cnn_comm* c, size_t count, int ch);
int comm_weights_ready(b200cnn_comm* c, cudaStream_t st);
// Make `st` wait for all communication enqueued so far (event join).
int comm_join(b200cnn_comm* c, cudaStream_t st);
// Gradient buffer this rank's backward must write for the step being enqueued, and the switch to
// the other half of the double buffer (called once per step, before the step is enqueued).
float* comm_current_grads(b200cnn_comm* c);
void comm_flip_parity(b200cnn_comm* c);
int comm_parity(const b200cnn_comm* c);

}  // namespace b200cnn
