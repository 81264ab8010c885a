// Internal interface of the NVLink communicator (see include/b200cnn_comm.h for the C ABI).
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

struct b200cnn_comm;

namespace b200cnn {

// Enqueue the fused all-reduce + 1/world + SGD for floats [offset, offset + count) of the flat
// parameter / gradient buffers.  The kernel runs on the communicator's own stream, ordered after
// everything enqueued on `after` so far (event fork), so it overlaps whatever `after` does next.
int comm_allreduce_sgd_slice(b200cnn_comm* c, size_t offset, size_t count, const float* lr_dev, cudaStream_t after);
// The same update enqueued on `st` itself, on a second flag channel (may overlap a slice on the communicator's stream).
int comm_allreduce_sgd_inline(b200cnn_comm* c, size_t offset, size_t count, const float* lr_dev, cudaStream_t st);
// Make `st` wait for all communication enqueued so far (event join).
int comm_join(b200cnn_comm* c, cudaStream_t st);
// Gradient buffer this rank's backward must write for the step being enqueued, and the switch to
// the other half of the double buffer (called once per step, before the step is enqueued).
float* comm_current_grads(b200cnn_comm* c);
void comm_flip_parity(b200cnn_comm* c);
int comm_parity(const b200cnn_comm* c);

}  // namespace b200cnn
