// One-hop push protocol of the NVLink communicator ("LL": low latency).
//
// The pull protocol of allreduce_sgd.cu costs two to three dependent NVLink trips per exchange: publish a flag with
// st.release.sys (a system-scope fence), spin on the peers' flags, then LOAD the peers' data across the link.  Here
// the producer pushes instead: every 8 bytes it writes into a peer's inbox are {4 bytes of payload, 4-byte tag}, two
// such pairs per 16-byte store, so data and "it has arrived" travel in the same packet.  The consumer polls its OWN
// memory until both tags of a packet equal the tag of the current call — no separate flag, no fence, no remote load:
// one one-way trip.  (8-byte stores are single-copy atomic; nothing wider is assumed.)
//
// Inboxes live in every rank's region behind the gradient buffers:
//     inbox[channel][buffer][source rank] = slot of `cap` bytes (payload capacity cap / 2)
// Tags are the per-channel call counter (1, 2, 3, ...; memory starts zeroed, 0 never matches).  The buffer is
// tag & 1: a rank can be at most one call ahead of a peer on a channel (it cannot finish call k before the peer has
// pushed its part of call k, and the peer pushes call k only after it has consumed call k-1), so two buffers are
// enough and no exit barrier exists.  Channels are independent call sequences (communicator stream, main stream,
// in-kernel tail) so that calls on different streams never share a counter.
//
// Replaces the per-sample host-memory MPI_Allreduce of /root/reference/cnnmpi.c:490.
#pragma once
#include <cstddef>
#include <cstdint>

#include "b200cnn_comm.h"

namespace b200cnn {
namespace ll {

constexpr int kChannels = 3;  // 0: communicator's own stream, 1: caller's stream (inline), 2: inside another kernel
constexpr int kMaxRanks = B200CNN_COMM_MAX_RANKS;
// Slot capacity in BYTES of inbox memory (payload = half of it).
__host__ __device__ constexpr size_t slot_bytes(int ch) { return ch == 0 ? (2u << 20) : ch == 1 ? (256u << 10) : (8u << 10); }
__host__ __device__ constexpr size_t channel_base(int ch) {
    size_t off = 0;
    for (int c = 0; c < ch; c++) off += 2 * kMaxRanks * slot_bytes(c);
    return off;
}
constexpr size_t kTotalBytes = channel_base(kChannels);

// Per-channel words in the communicator's local counter block (device memory of the owning rank only).
struct Counters {
    uint32_t calls;    // tag of the last completed call
    uint32_t done;     // blocks of the running call that have finished (reset by the last one)
    uint32_t wtag;     // tag of the last call whose result is written into this rank by its peers (two-shot)
    uint32_t wblocks;  // grid size of that call (= flags to wait for per peer)
};

struct Channel {
    int world = 1, rank = 0, ch = 0;
    char* region[kMaxRanks] = {};
    size_t inbox_off = 0;   // byte offset of the LL area inside a region
    Counters* ctr = nullptr;
    uint32_t* errflag = nullptr;
};

__device__ __forceinline__ uint4* slot_ptr(const Channel& c, int dst_rank, int buf, int src_rank) {
    return reinterpret_cast<uint4*>(c.region[dst_rank] + c.inbox_off + channel_base(c.ch) +
                                    size_t(buf * kMaxRanks + src_rank) * slot_bytes(c.ch));
}

__device__ __forceinline__ void st_packet(uint4* p, float a, float b, uint32_t tag) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(__float_as_uint(a)), "r"(tag), "r"(__float_as_uint(b)),
                 "r"(tag)
                 : "memory");
}
__device__ __forceinline__ uint4 ld_packet(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

// Four floats = two packets = one 32-byte sector of the destination's inbox, at float4 index i of the slot.
__device__ __forceinline__ void push4(uint4* slot, size_t i, const float4& v, uint32_t tag) {
    st_packet(slot + 2 * i, v.x, v.y, tag);
    st_packet(slot + 2 * i + 1, v.z, v.w, tag);
}

// Single floats (the in-kernel tail exchanges 160 of them): one 8-byte {value, tag} store per float.
__device__ __forceinline__ void push1(uint4* slot, size_t i, float v, uint32_t tag) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" ::"l"(reinterpret_cast<uint2*>(slot) + i), "r"(__float_as_uint(v)), "r"(tag)
                 : "memory");
}

constexpr long long kSpinTimeoutCycles = 4000000000ll;  // ~2 s

// Spin on the LOCAL inbox until float4 i of `slot` carries `tag`.  Returns false after ~2 s (peer never showed up).
__device__ __forceinline__ bool poll4(const uint4* slot, size_t i, uint32_t tag, float4& out) {
    long long t0 = 0;
    for (;;) {
        const uint4 a = ld_packet(slot + 2 * i), b = ld_packet(slot + 2 * i + 1);
        if (a.y == tag && a.w == tag && b.y == tag && b.w == tag) {
            out = make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x), __uint_as_float(b.z));
            return true;
        }
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > kSpinTimeoutCycles) return false;
    }
}

__device__ __forceinline__ bool poll1(const uint4* slot, size_t i, uint32_t tag, float& out) {
    long long t0 = 0;
    for (;;) {
        uint2 a;
        asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(a.x), "=r"(a.y) : "l"(reinterpret_cast<const uint2*>(slot) + i) : "memory");
        if (a.y == tag) {
            out = __uint_as_float(a.x);
            return true;
        }
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > kSpinTimeoutCycles) return false;
    }
}

}  // namespace ll
}  // namespace b200cnn
