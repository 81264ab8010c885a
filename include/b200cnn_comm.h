/* C ABI of the NVLink communicator (part of libb200cnn_cuda.so).
 *
 * What it replaces: the reference's only collective, a blocking host-memory
 *   MPI_Allreduce(MPI_IN_PLACE, ..., MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD)
 * issued six times per training sample on the wrong buffer, followed by a broken "averaging" loop
 * and a per-sample Layer_update (/root/reference/cnnmpi.c:487-499, CUDAMPI.c:412-421).  The
 * intended operation — sum the mini-batch weight/bias gradients over ranks, scale by 1/world,
 * SGD step — is ONE kernel here: every rank's kernel reads the peers' gradient buffers directly
 * over NVLink (P2P loads, or NVLS multimem instructions through the switch), and writes
 * w -= lr/world * sum.  No MPI, NCCL or separate elementwise kernel is involved.
 *
 * Memory model: each rank owns one "region" of device memory that every other rank has mapped
 * (CUDA IPC, CUDA VMM or torch symmetric memory — the library only needs the pointers):
 *     [ flags 16 KiB | params | grads[0] | grads[1] | push-protocol inboxes (36 MiB) ]
 * Gradients are double-buffered by step parity, so a rank may start the next step's backward
 * while slower peers are still reading the previous step's gradients, and no trailing barrier is
 * needed.  Cross-GPU ordering uses monotonically increasing sequence numbers in the flag words
 * (st.release.sys / ld.acquire.sys) or, in the push protocol, a tag inside every 8-byte packet;
 * nothing is ever reset.
 *
 * Ordering contract: all calls of one communicator that use the same push channel must be ordered
 * with respect to each other (same stream, or streams ordered by events) and made in the same order
 * on every rank.  The stand-alone calls below all use channel 0.
 */
#ifndef B200CNN_COMM_H
#define B200CNN_COMM_H

#include <stddef.h>
#include <stdint.h>

#include "b200cnn.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200cnn_comm b200cnn_comm;

enum {
    B200CNN_ALGO_AUTO = 0,
    B200CNN_ALGO_ONE_SHOT = 1, /* every rank reads all peers' full slice; best for small messages */
    B200CNN_ALGO_TWO_SHOT = 2, /* reduce-scatter + the owner writes UPDATED WEIGHTS to all peers */
    B200CNN_ALGO_NVLS = 3,     /* two-shot through the switch: multimem.ld_reduce + multimem.st */
    /* Push protocol (csrc/comm/ll.cuh): producers store {payload, tag} packets into the consumers' inboxes and the
     * consumers poll their own memory - one one-way NVLink trip, no flag exchange, no fence, no remote load. */
    B200CNN_ALGO_LL_ONE_SHOT = 4, /* everybody pushes the whole slice to everybody; slices up to an inbox slot */
    B200CNN_ALGO_LL_TWO_SHOT = 5  /* push shard s to rank s, the owner stores the UPDATED WEIGHTS into every peer */
};

#define B200CNN_COMM_MAX_RANKS 8
#define B200CNN_COMM_FLAG_BYTES 16384

/* Bytes of one rank's region for a flat parameter buffer of `flat_floats` floats. */
size_t b200cnn_comm_region_bytes(size_t flat_floats);
/* Offsets (bytes) of params / grads[parity] inside a region. */
size_t b200cnn_comm_params_offset(size_t flat_floats);
size_t b200cnn_comm_grads_offset(size_t flat_floats, int parity);

/* region_ptrs[r] = rank r's region as mapped in THIS process (region_ptrs[rank] is the local one).
 * multicast_ptr = NVLS multicast mapping of the same region, or NULL.  The regions must be zeroed
 * (at least the flag block) before any rank creates its communicator. */
b200cnn_comm* b200cnn_comm_create(int rank, int world, void* const* region_ptrs, void* multicast_ptr,
                                  size_t flat_floats);
void b200cnn_comm_destroy(b200cnn_comm* c);
int b200cnn_comm_rank(const b200cnn_comm* c);
int b200cnn_comm_world(const b200cnn_comm* c);
int b200cnn_comm_has_multicast(const b200cnn_comm* c);
float* b200cnn_comm_params(b200cnn_comm* c);
float* b200cnn_comm_grads(b200cnn_comm* c, int parity);
/* Algorithm used by the engine's bucket launches.  AUTO: push one-shot while (world-1) x bytes <= 448 KiB; NVLS from
 * 256 KiB when a multicast mapping exists and world >= 4; else push two-shot while a 1/world shard fits an inbox slot
 * (1 MiB) and the slice is <= 8 MiB; beyond that the pull kernels (two-shot, or NVLS from 8 MiB).  B200CNN_LL=0 in the
 * environment removes the push algorithms from AUTO. */
int b200cnn_comm_set_algo(b200cnn_comm* c, int algo);
/* 0 if no kernel of this communicator ever timed out waiting for a peer; the kernels give up after
 * ~2 s instead of hanging the GPU. */
int b200cnn_comm_error(b200cnn_comm* c);

/* Fused all-reduce(SUM) + 1/world + SGD over floats [offset, offset+count) (both multiples of 4):
 *     params[i] -= lr / world * sum_r grads_r[parity][i]      on every rank, bitwise identical.
 * Enqueued on `stream`; all ranks must call it with the same arguments in the same order. */
int b200cnn_comm_allreduce_sgd(b200cnn_comm* c, size_t offset, size_t count, float lr, int algo, int parity,
                               int max_blocks, b200cnn_stream stream);
/* Plain fused all-reduce(SUM) into grads[parity] of every rank (no update) — used by the bandwidth
 * sweep so the numbers are comparable with an NCCL all-reduce. */
int b200cnn_comm_allreduce(b200cnn_comm* c, size_t offset, size_t count, int algo, int parity, int max_blocks,
                           b200cnn_stream stream);
/* Copy rank `root`'s parameters into every rank's parameter buffer over NVLink (the reference never
 * synchronises its replicas: SURVEY D7). */
int b200cnn_comm_broadcast_params(b200cnn_comm* c, int root, b200cnn_stream stream);
/* Device-side barrier across ranks. */
int b200cnn_comm_barrier(b200cnn_comm* c, b200cnn_stream stream);
/* Enqueue the wait for the peers' halves of the last LL_TWO_SHOT updates (a stand-alone b200cnn_comm_allreduce_sgd
 * already includes it; the engine defers it to the start of the next step).  Call it, then synchronise, before
 * reading parameters on the host after training through an engine. */
int b200cnn_comm_wait_weights(b200cnn_comm* c, b200cnn_stream stream);

/* Let an engine run its parameters / gradients inside the communicator's region and launch the
 * fused update per layer, overlapped with the rest of backward. */
int b200cnn_engine_attach_comm(b200cnn_engine* e, b200cnn_comm* c);

/* Framework-free peer mapping: cudaMalloc'd (zeroed) regions shared through 64-byte CUDA IPC handles.
 * A rank allocates its region, publishes the handle (any host channel), and opens its peers'. */
void* b200cnn_ipc_alloc(size_t bytes);
int b200cnn_ipc_free(void* ptr);
int b200cnn_ipc_get_handle(void* ptr, unsigned char* out64);
void* b200cnn_ipc_open(const unsigned char* handle64);
int b200cnn_ipc_close(void* ptr);

/* Test hook: run `world` virtual ranks of the same kernels inside ONE launch on ONE GPU (the ranks
 * are the grid's y dimension), over `world` regions allocated on that GPU.  Exercises the flag
 * protocol and the reduction order without needing several GPUs. */
int b200cnn_comm_emulate_allreduce_sgd(int world, void* const* region_ptrs, size_t flat_floats, size_t offset,
                                       size_t count, float lr, int algo, int parity, int blocks, uint32_t* seq_scratch,
                                       b200cnn_stream stream);
/* Same, on flag slots [slot_base, slot_base + blocks): the engine runs a step's last bucket on slots 64.. so it can
 * overlap a bucket still running on slots 0.. from the communicator's own stream. */
int b200cnn_comm_emulate_allreduce_sgd_slot(int world, void* const* region_ptrs, size_t flat_floats, size_t offset,
                                            size_t count, float lr, int algo, int parity, int blocks, int slot_base,
                                            uint32_t* seq_scratch, b200cnn_stream stream);

/* Test hook for the push protocol, same idea: `world` virtual ranks in one launch.  scratch: zeroed device memory of
 * b200cnn_comm_emulate_ll_scratch_bytes(world) bytes, kept between calls (call counters + error word at its end).
 * update != 0: all-reduce + SGD; 0: all-reduce into grads[parity]. */
size_t b200cnn_comm_emulate_ll_scratch_bytes(int world);
int b200cnn_comm_emulate_ll(int world, void* const* region_ptrs, size_t flat_floats, size_t offset, size_t count, float lr,
                            int two_shot, int parity, int blocks, int channel, int update, void* scratch,
                            b200cnn_stream stream);

#ifdef __cplusplus
}
#endif
#endif
