// Device side of TailXchg (comm.cuh): a kernel that has just produced a few hundred floats (the weight gradient of
// the first convolution, in the block that finishes last) all-reduces them with its peers ITSELF instead of handing
// them to a separate collective kernel: compute + collective + update in one launch.  Push protocol of ll.cuh on
// channel 2: every float is stored into every peer's inbox as an 8-byte {value, tag} packet, then the CTA polls its
// own inbox — one one-way NVLink trip, no flag, no fence, no remote load.
#pragma once
#include <cstdint>

#include "comm.cuh"

namespace b200cnn {
namespace xchg {

// Called by EVERY thread of ONE CTA per rank.  Thread t < n contributes `mine` (element t of the slice) and receives in
// `sum` the total over ranks, added in rank order (bitwise identical on every rank).  Returns false for a thread
// whose peer never showed up (~2 s; the communicator's error flag is raised and the caller must skip its update).
__device__ __forceinline__ bool allreduce_small(const TailXchg& x, int t, int n, float mine, float& sum) {
    const ll::Channel& c = x.chan;
    const uint32_t tag = c.ctr->calls + 1;
    const int buf = int(tag & 1u);
    bool ok = true;
    if (t < n) {
        for (int k = 0; k < c.world; k++)
            if (k != c.rank) ll::push1(ll::slot_ptr(c, k, buf, c.rank), size_t(t), mine, tag);
        float acc = 0.f;
        for (int k = 0; k < c.world && ok; k++) {
            float v = mine;
            if (k != c.rank) ok = ll::poll1(ll::slot_ptr(c, c.rank, buf, k), size_t(t), tag, v);
            acc = k == 0 ? v : acc + v;
        }
        sum = acc;
        if (!ok) atomicExch(c.errflag, 1u);
    }
    __syncthreads();  // every thread has read the call counter
    if (threadIdx.x == 0) c.ctr->calls = tag;
    return ok;
}

}  // namespace xchg
}  // namespace b200cnn
