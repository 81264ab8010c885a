// Device side of TailXchg (comm.cuh): the flag protocol of allreduce_sgd.cu (per-slot sequence numbers that only
// grow, st.release.sys after a CTA barrier, ld.acquire.sys polling, bounded wait) for a kernel that exchanges a few
// hundred floats with its peers in its own last block instead of handing them to a separate collective kernel.
#pragma once
#include <cstdint>

#include "comm.cuh"

namespace b200cnn {
namespace xchg {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_peer(const float* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
// Same layout as flag_ptr() in allreduce_sgd.cu: [slot][phase][source rank] words at the start of a region.
__device__ __forceinline__ uint32_t* flag_ptr(char* region, int slot, int phase, int src) {
    return reinterpret_cast<uint32_t*>(region) + (slot * 2 + phase) * B200CNN_COMM_MAX_RANKS + src;
}

// Every thread of the CTA calls this after the CTA's results are stored in this rank's gradient buffer.  On return
// every peer's results of the same step are stored too (and visible).  Returns the sequence number to be written
// back to *x.seq by ONE thread once the CTA is done with the exchange.
__device__ __forceinline__ uint32_t all_ranks_ready(const TailXchg& x) {
    const uint32_t seq = *x.seq + 1;
    __syncthreads();
    if (int(threadIdx.x) < x.world && int(threadIdx.x) != x.rank) {
        const int peer = threadIdx.x;
        st_release_sys(flag_ptr(x.region[peer], x.slot, 0, x.rank), seq);
        const uint32_t* mine = flag_ptr(x.region[x.rank], x.slot, 0, peer);
        const long long t0 = clock64();
        while (int32_t(ld_acquire_sys(mine) - seq) < 0) {
            if (clock64() - t0 > 4000000000ll) {  // ~2 s: raise the communicator's error flag instead of hanging
                atomicExch(x.errflag, 1u);
                break;
            }
        }
    }
    __syncthreads();
    return seq;
}

}  // namespace xchg
}  // namespace b200cnn
