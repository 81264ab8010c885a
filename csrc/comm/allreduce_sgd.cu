// Fused gradient all-reduce + 1/world scaling + SGD update over NVLink 5 / NVSwitch (K8), the weight
// broadcast (K9) and a device-side barrier.  See include/b200cnn_comm.h for the contract and the
// reference code this replaces (/root/reference/cnnmpi.c:487-499 + cnn.c:303-311).
//
// Three variants of the same operation:
//   one-shot  every rank loads the slice from all W ranks (its own included) in rank order, adds,
//             updates its own weights.  (W-1)*N bytes in per GPU, one flag exchange.  Small slices.
//   two-shot  rank r owns 1/W of the slice: loads that shard from all ranks, adds, computes the
//             UPDATED WEIGHTS and stores them into every rank's parameter buffer.  2*(W-1)/W*N bytes
//             per direction per GPU, two flag exchanges.  Replicas are identical by construction.
//   NVLS      two-shot where the switch does the adding and the fan-out:
//             multimem.ld_reduce.add.v4.f32 on the multicast address, multimem.st of the new weights.
//
// Ordering: flag words carry a per-block sequence number that only ever grows; a rank publishes with
// st.release.sys after a CTA barrier and waits with ld.acquire.sys.  Waits give up after ~2 s and
// raise an error flag rather than hanging the GPU.
#include <cstdlib>
#include <new>

#include "../cuda/common.cuh"
#include "b200cnn_comm.h"
#include "comm.cuh"
#include "ll.cuh"

namespace b200cnn {

constexpr int kMaxRanks = B200CNN_COMM_MAX_RANKS;
constexpr int kMaxBlocks = 128;
constexpr int kCommThreads = 512;
constexpr long long kSpinTimeoutCycles = 4000000000ll;  // ~2 s at 1.9 GHz

static_assert(kMaxBlocks * 2 * kMaxRanks * sizeof(uint32_t) <= B200CNN_COMM_FLAG_BYTES, "flag block too small");

struct CommView {
    int rank, world, emulate;
    char* region[kMaxRanks];
    char* mc_region;
    size_t params_off, grads_off[2];
    uint32_t* seq;     // local sequence counters [emulate ? world : 1][kMaxBlocks]
    uint32_t* errflag; // local
    size_t ll_off;     // byte offset of the push-protocol inboxes inside a region (ll.cuh)
    ll::Counters* ll_ctr;  // local [emulate ? world : 1][ll::kChannels]
};

constexpr int kLLMaxBlocks = 64;
// "Weights have landed" flags of the two-shot push algorithm: [channel 0..1][source rank][block] words in the second
// half of the flag block.
constexpr size_t kWFlagOffset = 8192;
static_assert(kWFlagOffset + 2 * kMaxRanks * kLLMaxBlocks * sizeof(uint32_t) <= B200CNN_COMM_FLAG_BYTES, "flag block too small");

enum Op { kOpSgd = 0, kOpReduceOnly = 1, kOpBroadcast = 2, kOpBarrier = 3 };

namespace {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Peer data: coherent at system scope, never served from a stale L1 line.
__device__ __forceinline__ float4 ld_peer(const float4* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ void st_peer(float4* p, const float4& v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float4* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float4* mc, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

__device__ __forceinline__ uint32_t* flag_ptr(char* region, int block, int phase, int src) {
    return reinterpret_cast<uint32_t*>(region) + (block * 2 + phase) * kMaxRanks + src;
}

// Cross-rank barrier between the blocks with this blockIdx.x on every rank.  All earlier global
// writes of the CTA are published (CTA barrier, then a cumulative release at system scope).
// Returns true if a peer never showed up (~2 s): the caller must then skip its update and stores — reducing
// incomplete data would silently corrupt the replicas; the error flag is what the host polls.
__device__ bool signal_and_wait(const CommView& v, int rank, int phase, uint32_t seq, int slot) {
    __syncthreads();
    int timed_out = 0;
    if (threadIdx.x < v.world && int(threadIdx.x) != rank) {
        const int peer = threadIdx.x;
        st_release_sys(flag_ptr(v.region[peer], slot, phase, rank), seq);
        const uint32_t* mine = flag_ptr(v.region[rank], slot, phase, peer);
        const long long t0 = clock64();
        // (Measured: polling with ld.acquire.sys is ~2 us faster per exchange than relaxed polling followed by
        //  fence.acq_rel.sys, whose system-scope fence alone costs more than the whole spin.)
        while (int32_t(ld_acquire_sys(mine) - seq) < 0) {
            if (clock64() - t0 > kSpinTimeoutCycles) {
                atomicExch(v.errflag, 1u);
                timed_out = 1;
                break;
            }
        }
    }
    return __syncthreads_or(timed_out) != 0;
}

__device__ __forceinline__ float4 add4(float4 a, const float4& b) {
    a.x += b.x;
    a.y += b.y;
    a.z += b.z;
    a.w += b.w;
    return a;
}

// Sum of element i over ranks 0..W-1 in that fixed order (identical on every rank).
template <int W>
__device__ __forceinline__ float4 reduce_ranks(const float4* const* g, int world, size_t i) {
    float4 vals[W > 0 ? W : 1];
    if (W > 0) {
#pragma unroll
        for (int k = 0; k < W; k++) vals[k] = ld_peer(g[k] + i);  // all loads in flight before the adds
        float4 acc = vals[0];
#pragma unroll
        for (int k = 1; k < W; k++) acc = add4(acc, vals[k]);
        return acc;
    } else {
        float4 acc = ld_peer(g[0] + i);
        for (int k = 1; k < world; k++) acc = add4(acc, ld_peer(g[k] + i));
        return acc;
    }
}

// Elements per thread per trip: a trip costs one NVLink round trip whatever it carries, so small worlds put
// several elements' loads in flight at once (U x W = 8 peer loads per thread per trip).
template <int W>
struct Unroll {
    static constexpr int value = (W == 1 || W == 2) ? 4 : (W == 4 ? 2 : 1);
};

// s[u] = sum over ranks (rank order) of element i + u * stride, for the u with i + u * stride < end.
template <int W, int U>
__device__ __forceinline__ void reduce_ranks_batch(const float4* const* g, int world, size_t i, size_t stride, size_t end,
                                                   float4 (&s)[U]) {
    if (W > 0 && U > 1) {
        float4 vals[U][W > 0 ? W : 1];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const size_t idx = i + size_t(u) * stride;
            if (idx < end) {
#pragma unroll
                for (int k = 0; k < W; k++) vals[u][k] = ld_peer(g[k] + idx);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (i + size_t(u) * stride < end) {
                float4 acc = vals[u][0];
#pragma unroll
                for (int k = 1; k < W; k++) acc = add4(acc, vals[u][k]);
                s[u] = acc;
            }
        }
    } else {
#pragma unroll
        for (int u = 0; u < U; u++)
            if (i + size_t(u) * stride < end) s[u] = reduce_ranks<W>(g, world, i + size_t(u) * stride);
    }
}

template <int W>
__global__ void __launch_bounds__(kCommThreads)
allreduce_kernel(CommView v, size_t off4, size_t n4, const float* __restrict__ lr_dev, float lr_imm, int algo,
                 int parity, int op, int root, int block_base) {
    const int rank = v.emulate ? int(blockIdx.y) : v.rank;
    const int world = v.world;
    // block_base picks the flag words / sequence counters: two kernels of one communicator may run concurrently
    // (different streams) when their [block_base, block_base + gridDim.x) ranges are disjoint.
    const int slot = int(blockIdx.x) + block_base;
    uint32_t* seq_slot = v.seq + (v.emulate ? rank : 0) * kMaxBlocks + slot;
    const uint32_t seq = *seq_slot + 1;

    if (signal_and_wait(v, rank, 0, seq, slot)) return;  // every rank's inputs for this call are complete (or: give up)

    const float4* g[kMaxRanks];
#pragma unroll
    for (int k = 0; k < kMaxRanks; k++)
        g[k] = reinterpret_cast<const float4*>(v.region[k < world ? k : 0] + v.grads_off[parity]) + off4;
    float4* w_local = reinterpret_cast<float4*>(v.region[rank] + v.params_off) + off4;
    float4* g_local = reinterpret_cast<float4*>(v.region[rank] + v.grads_off[parity]) + off4;
    const float lr = (lr_dev ? *lr_dev : lr_imm) / float(world);
    const size_t tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x, nthreads = size_t(gridDim.x) * blockDim.x;

    bool need_exit_barrier = false;
    if (op == kOpBarrier) {
        // nothing to move
    } else if (op == kOpBroadcast) {
        const float4* src = reinterpret_cast<const float4*>(v.region[root] + v.params_off) + off4;
        if (rank != root)
            for (size_t i = tid; i < n4; i += nthreads) w_local[i] = ld_peer(src + i);
        need_exit_barrier = true;  // root must not change its parameters before everyone has read them
    } else if (algo == B200CNN_ALGO_ONE_SHOT) {
        if (op == kOpSgd) {
            constexpr int U = Unroll<W>::value;
            for (size_t i = tid; i < n4; i += nthreads * U) {
                float4 s[U], w[U];
#pragma unroll
                for (int u = 0; u < U; u++)
                    if (i + u * nthreads < n4) w[u] = w_local[i + u * nthreads];
                reduce_ranks_batch<W, U>(g, world, i, nthreads, n4, s);
#pragma unroll
                for (int u = 0; u < U; u++) {
                    if (i + u * nthreads < n4) {
                        w[u].x -= lr * s[u].x;
                        w[u].y -= lr * s[u].y;
                        w[u].z -= lr * s[u].z;
                        w[u].w -= lr * s[u].w;
                        w_local[i + u * nthreads] = w[u];
                    }
                }
            }
        } else {
            // In-place all-reduce: peers are still reading this rank's slice, so results may only
            // be written after everyone has finished reading.
            // (Each thread keeps its results in registers across the barrier when they fit; otherwise
            //  this falls back to two-shot, which the host selects for large slices anyway.)
            float4 keep[4];
            int nkeep = 0;
            for (size_t i = tid; i < n4 && nkeep < 4; i += nthreads) keep[nkeep++] = reduce_ranks<W>(g, world, i);
            if (signal_and_wait(v, rank, 1, seq, slot)) return;
            nkeep = 0;
            for (size_t i = tid; i < n4 && nkeep < 4; i += nthreads) g_local[i] = keep[nkeep++];
        }
    } else {
        // two-shot / NVLS: this rank owns [lo, hi) of the slice
        const size_t lo = n4 * size_t(rank) / world, hi = n4 * size_t(rank + 1) / world;
        if (algo == B200CNN_ALGO_NVLS) {
            const float4* mg = reinterpret_cast<const float4*>(v.mc_region + v.grads_off[parity]) + off4;
            float4* mw = reinterpret_cast<float4*>(v.mc_region + v.params_off) + off4;
            float4* mgw = reinterpret_cast<float4*>(v.mc_region + v.grads_off[parity]) + off4;
            for (size_t i = lo + tid; i < hi; i += nthreads) {
                const float4 s = multimem_ld_reduce_add(mg + i);
                if (op == kOpSgd) {
                    float4 w = w_local[i];
                    w.x -= lr * s.x;
                    w.y -= lr * s.y;
                    w.z -= lr * s.z;
                    w.w -= lr * s.w;
                    multimem_st(mw + i, w);
                } else {
                    multimem_st(mgw + i, s);
                }
            }
        } else {
            constexpr int U = Unroll<W>::value;
            const size_t dst_off = (op == kOpSgd) ? v.params_off : v.grads_off[parity];
            for (size_t i = lo + tid; i < hi; i += nthreads * U) {
                float4 s[U], out[U];
                if (op == kOpSgd) {
#pragma unroll
                    for (int u = 0; u < U; u++)
                        if (i + u * nthreads < hi) out[u] = w_local[i + u * nthreads];
                }
                reduce_ranks_batch<W, U>(g, world, i, nthreads, hi, s);
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const size_t idx = i + u * nthreads;
                    if (idx >= hi) continue;
                    if (op == kOpSgd) {
                        out[u].x -= lr * s[u].x;
                        out[u].y -= lr * s[u].y;
                        out[u].z -= lr * s[u].z;
                        out[u].w -= lr * s[u].w;
                    } else {
                        out[u] = s[u];
                    }
#pragma unroll
                    for (int k = 0; k < kMaxRanks; k++)
                        if (k < world) st_peer(reinterpret_cast<float4*>(v.region[k] + dst_off) + off4 + idx, out[u]);
                }
            }
        }
        need_exit_barrier = true;  // all shards must have landed everywhere before anyone moves on
    }
    if (need_exit_barrier) signal_and_wait(v, rank, 1, seq, slot);
    if (threadIdx.x == 0) *seq_slot = seq;
}

// ---- push protocol (ll.cuh): one-hop all-reduce + SGD -------------------------------------------------------------
__device__ __forceinline__ uint32_t* wflag_ptr(char* region, int ch, int src, int block) {
    return reinterpret_cast<uint32_t*>(region + kWFlagOffset) + (ch * kMaxRanks + src) * kLLMaxBlocks + block;
}

__device__ __forceinline__ ll::Channel make_channel(const CommView& v, int rank, int ch) {
    ll::Channel c;
    c.world = v.world;
    c.rank = rank;
    c.ch = ch;
#pragma unroll
    for (int k = 0; k < kMaxRanks; k++) c.region[k] = v.region[k];
    c.inbox_off = v.ll_off;
    c.ctr = v.ll_ctr + (v.emulate ? rank : 0) * ll::kChannels + ch;
    c.errflag = v.errflag;
    return c;
}

// one-shot (two_shot == 0): every rank pushes its whole slice into every peer's inbox, then adds what arrived in its
//   own inbox to its own gradient in rank order and updates its own weights.  One one-way NVLink trip, no barrier.
// two-shot: rank r owns shard r.  Every rank pushes shard s of its gradient to rank s (reduce-scatter, one trip);
//   the owner adds in rank order, computes the UPDATED WEIGHTS and stores them into every rank's parameter buffer
//   (second trip), then raises one flag per block in every peer.  Nobody waits for those flags here: the consumer of
//   the weights does, one step later (ll_wait_weights_kernel), so the second trip is off the critical path.
template <int W>
__global__ void __launch_bounds__(kCommThreads)
ll_allreduce_kernel(CommView v, size_t off4, size_t n4, const float* __restrict__ lr_dev, float lr_imm, int two_shot,
                    int parity, int op, int ch) {
    const int rank = v.emulate ? int(blockIdx.y) : v.rank;
    const int world = W > 0 ? W : v.world;
    const ll::Channel c = make_channel(v, rank, ch);
    const uint32_t tag = c.ctr->calls + 1;  // bumped by the block that finishes last, i.e. after every block has read it
    const int buf = int(tag & 1u);
    float4* w_local = reinterpret_cast<float4*>(v.region[rank] + v.params_off) + off4;
    float4* g_local = reinterpret_cast<float4*>(v.region[rank] + v.grads_off[parity]) + off4;
    const float lr = (lr_dev ? *lr_dev : lr_imm) / float(world);
    const size_t tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x, nthreads = size_t(gridDim.x) * blockDim.x;
    bool ok = true;

    // Every loop below first puts all of an iteration's loads in flight (its own gradients / weights and a first look
    // at every peer's packets) and only then consumes them: the kernel moves a few hundred KB per rank and was bound
    // by one dependent memory latency per element before (22 us for 1.4 MB at 2 ranks).
    constexpr int WW = W > 0 ? W : kMaxRanks;
    constexpr int U = Unroll<W>::value;  // elements per thread per trip: U x (W-1) packets in flight
    auto reduce_batch = [&](size_t i, size_t stride, size_t end, size_t slot_base, float4 (&acc)[U]) {
        uint4 pa[U][WW], pb[U][WW];
        float4 own[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const size_t idx = i + size_t(u) * stride;
            if (idx >= end) continue;
            own[u] = g_local[idx];
#pragma unroll
            for (int k = 0; k < WW; k++)
                if (k < world && k != rank) {
                    const uint4* slot = ll::slot_ptr(c, rank, buf, k);
                    pa[u][k] = ll::ld_packet(slot + 2 * (idx - slot_base));
                    pb[u][k] = ll::ld_packet(slot + 2 * (idx - slot_base) + 1);
                }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const size_t idx = i + size_t(u) * stride;
            if (idx >= end) continue;
            float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int k = 0; k < WW; k++) {
                if (k >= world) continue;
                float4 t4 = own[u];
                if (k != rank) {
                    const uint4 x = pa[u][k], y = pb[u][k];
                    if (x.y == tag && x.w == tag && y.y == tag && y.w == tag)
                        t4 = make_float4(__uint_as_float(x.x), __uint_as_float(x.z), __uint_as_float(y.x), __uint_as_float(y.z));
                    else
                        ok = ok && ll::poll4(ll::slot_ptr(c, rank, buf, k), idx - slot_base, tag, t4);  // not there yet: spin
                }
                sum = k == 0 ? t4 : add4(sum, t4);  // rank order: bitwise identical on every rank
            }
            acc[u] = sum;
        }
    };

    if (!two_shot) {
        for (size_t i = tid; i < n4; i += nthreads * 4) {
            float4 mine[4];
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (i + u * nthreads < n4) mine[u] = g_local[i + u * nthreads];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                if (i + u * nthreads >= n4) continue;
#pragma unroll
                for (int k = 0; k < WW; k++)
                    if (k < world && k != rank) ll::push4(ll::slot_ptr(c, k, buf, rank), i + u * nthreads, mine[u], tag);
            }
        }
        for (size_t i = tid; i < n4 && ok; i += nthreads * U) {
            float4 acc[U], w[U];
            if (op == kOpSgd) {
#pragma unroll
                for (int u = 0; u < U; u++)
                    if (i + u * nthreads < n4) w[u] = w_local[i + u * nthreads];
            }
            reduce_batch(i, nthreads, n4, 0, acc);
            if (!ok) break;
#pragma unroll
            for (int u = 0; u < U; u++) {
                const size_t idx = i + u * nthreads;
                if (idx >= n4) continue;
                if (op == kOpSgd) {
                    w[u].x -= lr * acc[u].x;
                    w[u].y -= lr * acc[u].y;
                    w[u].z -= lr * acc[u].z;
                    w[u].w -= lr * acc[u].w;
                    w_local[idx] = w[u];
                } else {
                    g_local[idx] = acc[u];
                }
            }
        }
    } else {
        // reduce-scatter: my part of shard s goes to rank s; one element of every peer's shard per trip
        size_t longest = 0;
#pragma unroll
        for (int k = 0; k < WW; k++)
            if (k < world) longest = max(longest, n4 * size_t(k + 1) / world - n4 * size_t(k) / world);
        for (size_t base = tid; base < longest; base += nthreads) {
            float4 mine[WW];
#pragma unroll
            for (int k = 0; k < WW; k++) {
                if (k >= world || k == rank) continue;
                const size_t lo = n4 * size_t(k) / world, hi = n4 * size_t(k + 1) / world;
                if (lo + base < hi) mine[k] = g_local[lo + base];
            }
#pragma unroll
            for (int k = 0; k < WW; k++) {
                if (k >= world || k == rank) continue;
                const size_t lo = n4 * size_t(k) / world, hi = n4 * size_t(k + 1) / world;
                if (lo + base < hi) ll::push4(ll::slot_ptr(c, k, buf, rank), base, mine[k], tag);
            }
        }
        const size_t lo = n4 * size_t(rank) / world, hi = n4 * size_t(rank + 1) / world;
        const size_t dst_off = (op == kOpSgd) ? v.params_off : v.grads_off[parity];
        for (size_t i = lo + tid; i < hi && ok; i += nthreads * U) {
            float4 acc[U], out[U];
            if (op == kOpSgd) {
#pragma unroll
                for (int u = 0; u < U; u++)
                    if (i + u * nthreads < hi) out[u] = w_local[i + u * nthreads];
            }
            reduce_batch(i, nthreads, hi, lo, acc);
            if (!ok) break;
#pragma unroll
            for (int u = 0; u < U; u++) {
                const size_t idx = i + u * nthreads;
                if (idx >= hi) continue;
                if (op == kOpSgd) {
                    out[u].x -= lr * acc[u].x;
                    out[u].y -= lr * acc[u].y;
                    out[u].z -= lr * acc[u].z;
                    out[u].w -= lr * acc[u].w;
                } else {
                    out[u] = acc[u];
                }
#pragma unroll
                for (int k = 0; k < WW; k++)
                    if (k < world) st_peer(reinterpret_cast<float4*>(v.region[k] + dst_off) + off4 + idx, out[u]);
            }
        }
    }
    if (!ok) atomicExch(v.errflag, 1u);
    __syncthreads();  // the block's stores are issued; one release per peer publishes them
    if (two_shot && threadIdx.x < world && int(threadIdx.x) != rank)
        st_release_sys(wflag_ptr(v.region[threadIdx.x], ch, rank, int(blockIdx.x)), tag);
    if (threadIdx.x == 0) {
        const uint32_t prev = atomicAdd(&c.ctr->done, 1u);
        if (prev == gridDim.x - 1) {  // last block of this rank: every block has read `calls`
            c.ctr->done = 0;
            c.ctr->calls = tag;
            if (two_shot) {
                c.ctr->wtag = tag;
                c.ctr->wblocks = gridDim.x;
            }
        }
    }
}

// Second half of the two-shot push algorithm, run by the CONSUMER of the result (the next step's forward, or the
// host before it reads the parameters): wait until every peer's blocks have raised their flag for the last two-shot
// call on channel `ch`.  One block.
__global__ void __launch_bounds__(kCommThreads) ll_wait_weights_kernel(CommView v, int ch) {
    const int rank = v.emulate ? int(blockIdx.y) : v.rank;
    const ll::Channel c = make_channel(v, rank, ch);
    const uint32_t wtag = c.ctr->wtag;
    const int nb = int(c.ctr->wblocks);
    if (wtag == 0) return;
    for (int idx = threadIdx.x; idx < v.world * nb; idx += blockDim.x) {
        const int src = idx / nb, b = idx % nb;
        if (src == rank) continue;
        const uint32_t* f = wflag_ptr(v.region[rank], ch, src, b);
        const long long t0 = clock64();
        while (int32_t(ld_acquire_sys(f) - wtag) < 0) {
            if (clock64() - t0 > kSpinTimeoutCycles) {
                atomicExch(v.errflag, 1u);
                break;
            }
        }
    }
}

int pick_blocks(size_t n4, int max_blocks, int per_thread = 4) {
    // ~4 float4 per thread per block before adding blocks: launch + flag latency dominates small slices.
    // (per_thread = 1 for a slice on the critical path: every extra trip of the loop is one more NVLink round trip,
    //  while extra blocks exchange their flags in parallel.)
    size_t b = (n4 + size_t(kCommThreads) * per_thread - 1) / (size_t(kCommThreads) * per_thread);
    if (max_blocks <= 0 || max_blocks > kMaxBlocks) max_blocks = kMaxBlocks;
    if (b < 1) b = 1;
    if (b > size_t(max_blocks)) b = size_t(max_blocks);
    return int(b);
}

}  // namespace
}  // namespace b200cnn

using namespace b200cnn;

struct b200cnn_comm {
    CommView view{};
    size_t flat_floats = 0;
    int algo = B200CNN_ALGO_AUTO;
    int parity = 0;
    uint32_t* local = nullptr;  // seq counters + error flag + push-protocol counters
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_wait = nullptr;
    bool forked = false;
    bool wait_pending = false;  // a wait kernel has been enqueued on `stream` and `ev_wait` recorded behind it
};

namespace {

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

bool ll_enabled() {
    static const bool on = [] {
        const char* v = std::getenv("B200CNN_LL");  // B200CNN_LL=0: pull protocol only (A/B timing)
        return !(v && v[0] == '0');
    }();
    return on;
}

int resolve_algo(const b200cnn_comm* c, int algo, size_t count, int op, int ch) {
    if (algo == B200CNN_ALGO_AUTO) algo = c->algo;
    if (c->view.world == 1) return B200CNN_ALGO_ONE_SHOT;
    const size_t payload = ll::slot_bytes(ch) / 2;  // bytes of payload one source may push into one inbox slot
    const size_t shard = (count / 4 + c->view.world - 1) / c->view.world * 16 + 16;
    if (algo == B200CNN_ALGO_AUTO && ll_enabled()) {
        // Push protocol wherever the slice fits an inbox: one-shot while every rank can afford to send the whole
        // slice to everybody ((W-1) x 2 x bytes over its links), two-shot while a shard fits a slot.
        // Measured (profiles/allreduce_sweep_w8.json, 8 ranks): push one-shot 9.4 / 12.6 us at 4 / 64 KB against 11.0 /
        // 23.0 for the pull kernel; from 256 KB the switch reduction wins where the box has it (NVLS 16.5 / 20.0 us at
        // 256 KB / 1.44 MB against 19.6 / 28.0 for push two-shot), so push two-shot is the choice only without multicast.
        const size_t bytes = count * sizeof(float);
        if (bytes <= payload && bytes * size_t(c->view.world - 1) <= (448u << 10)) return B200CNN_ALGO_LL_ONE_SHOT;
        if (c->view.mc_region && c->view.world >= 4 && bytes >= (256u << 10)) return B200CNN_ALGO_NVLS;
        if (shard <= payload && bytes <= (8u << 20)) return B200CNN_ALGO_LL_TWO_SHOT;
    }
    if (algo == B200CNN_ALGO_LL_ONE_SHOT && count * sizeof(float) > payload) algo = B200CNN_ALGO_LL_TWO_SHOT;
    if (algo == B200CNN_ALGO_LL_TWO_SHOT && shard > payload) algo = B200CNN_ALGO_AUTO;  // too big for the inboxes
    if (algo == B200CNN_ALGO_LL_ONE_SHOT || algo == B200CNN_ALGO_LL_TWO_SHOT) return algo;
    if (algo == B200CNN_ALGO_AUTO) {
        // Measured crossovers on 8xB200 (profiles/allreduce_sweep_w*.json and in-step runs): one-shot wins while
        // (W-1) x bytes of peer loads cost less than a second flag exchange; NVLS wins once the message is large
        // enough for the switch reduction to matter; two-shot in between.
        const size_t bytes = count * sizeof(float);
        const size_t one_shot_max = c->view.world <= 2 ? (4u << 20) : c->view.world <= 4 ? (256u << 10) : (8u << 10);
        if (bytes <= one_shot_max) algo = B200CNN_ALGO_ONE_SHOT;
        else if (bytes >= (8u << 20) && c->view.mc_region) algo = B200CNN_ALGO_NVLS;
        else algo = B200CNN_ALGO_TWO_SHOT;
    }
    if (algo == B200CNN_ALGO_NVLS && !c->view.mc_region) algo = B200CNN_ALGO_TWO_SHOT;
    if (algo == B200CNN_ALGO_AUTO) algo = B200CNN_ALGO_TWO_SHOT;
    // The in-place one-shot all-reduce keeps at most 4 float4 per thread in registers.
    if (op == kOpReduceOnly && algo == B200CNN_ALGO_ONE_SHOT &&
        count / 4 > size_t(kMaxBlocks) * kCommThreads * 4)
        algo = B200CNN_ALGO_TWO_SHOT;
    return algo;
}

int launch(const CommView& v, int ranks_y, size_t off4, size_t n4, const float* lr_dev, float lr, int algo, int parity,
           int op, int root, int blocks, cudaStream_t st, int block_base = 0) {
    if (block_base + blocks > kMaxBlocks) return set_error(kInvalidValue, "comm: %d blocks from slot %d exceed the %d flag slots", blocks, block_base, kMaxBlocks);
    const dim3 grid(blocks, ranks_y);
#define B200CNN_LAUNCH_W(W) allreduce_kernel<W><<<grid, kCommThreads, 0, st>>>(v, off4, n4, lr_dev, lr, algo, parity, op, root, block_base)
    switch (v.world) {
        case 1: B200CNN_LAUNCH_W(1); break;
        case 2: B200CNN_LAUNCH_W(2); break;
        case 4: B200CNN_LAUNCH_W(4); break;
        case 8: B200CNN_LAUNCH_W(8); break;
        default: B200CNN_LAUNCH_W(0); break;
    }
#undef B200CNN_LAUNCH_W
    B200CNN_CHECK_LAUNCH("allreduce_sgd");
    return 0;
}

int launch_ll(const CommView& v, int ranks_y, size_t off4, size_t n4, const float* lr_dev, float lr, int two_shot, int parity,
              int op, int ch, int blocks, cudaStream_t st) {
    const dim3 grid(blocks, ranks_y);
#define B200CNN_LAUNCH_LL(W) ll_allreduce_kernel<W><<<grid, kCommThreads, 0, st>>>(v, off4, n4, lr_dev, lr, two_shot, parity, op, ch)
    switch (v.world) {
        case 2: B200CNN_LAUNCH_LL(2); break;
        case 4: B200CNN_LAUNCH_LL(4); break;
        case 8: B200CNN_LAUNCH_LL(8); break;
        default: B200CNN_LAUNCH_LL(0); break;
    }
#undef B200CNN_LAUNCH_LL
    B200CNN_CHECK_LAUNCH("ll_allreduce_sgd");
    return 0;
}

int launch_ll_wait(const CommView& v, int ranks_y, int ch, cudaStream_t st) {
    ll_wait_weights_kernel<<<dim3(1, ranks_y), kCommThreads, 0, st>>>(v, ch);
    B200CNN_CHECK_LAUNCH("ll_wait_weights");
    return 0;
}

// Grid of a push-protocol call: the push loop is store-issue bound, so a big slice gets all the blocks it is allowed;
// a small one gets one float4 per thread (every extra trip of the poll loop is exposed latency).
int pick_ll_blocks(size_t n4, int max_blocks, bool two_shot) {
    if (max_blocks <= 0 || max_blocks > kLLMaxBlocks) max_blocks = two_shot ? 32 : 16;
    size_t b = (n4 + kCommThreads - 1) / kCommThreads;
    if (two_shot) b = (b + 3) / 4;
    if (b < 1) b = 1;
    if (b > size_t(max_blocks)) b = size_t(max_blocks);
    return int(b);
}

// ch: push-protocol channel (ll.cuh) = which stream family this call is ordered on.  defer_wait: a two-shot push call
// returns once THIS rank's shard is stored everywhere; the wait for the peers' shards is left to comm_wait_weights().
int enqueue(b200cnn_comm* c, size_t offset, size_t count, const float* lr_dev, float lr, int algo, int parity, int op,
            int root, int max_blocks, cudaStream_t st, int block_base = 0, int per_thread = 4, int ch = 0,
            bool defer_wait = false) {
    if ((offset | count) & 3) return set_error(kInvalidValue, "comm: offset and count must be multiples of 4 floats");
    if (offset + count > c->flat_floats) return set_error(kInvalidValue, "comm: slice [%zu, %zu) outside the %zu-float buffer", offset, offset + count, c->flat_floats);
    if (parity != 0 && parity != 1) return set_error(kInvalidValue, "comm: parity must be 0 or 1");
    algo = (op == kOpSgd || op == kOpReduceOnly) ? resolve_algo(c, algo, count, op, ch) : B200CNN_ALGO_ONE_SHOT;
    const size_t n4 = count / 4;
    if (algo == B200CNN_ALGO_LL_ONE_SHOT || algo == B200CNN_ALGO_LL_TWO_SHOT) {
        const bool two = algo == B200CNN_ALGO_LL_TWO_SHOT;
        if (int rc = launch_ll(c->view, 1, offset / 4, n4, lr_dev, lr, two ? 1 : 0, parity, op, ch, pick_ll_blocks(n4, max_blocks, two), st)) return rc;
        if (two && !defer_wait) return launch_ll_wait(c->view, 1, ch, st);
        return 0;
    }
    int blocks = pick_blocks(n4, max_blocks, per_thread);
    if (op == kOpReduceOnly && algo == B200CNN_ALGO_ONE_SHOT) {
        // registers hold the results across the barrier: make sure 4 per thread is enough
        const size_t need = (n4 + size_t(kCommThreads) * 4 - 1) / (size_t(kCommThreads) * 4);
        if (size_t(blocks) < need) blocks = int(need);
    }
    return launch(c->view, 1, offset / 4, n4, lr_dev, lr, algo, parity, op, root, blocks, st, block_base);
}

}  // namespace

extern "C" size_t b200cnn_comm_params_offset(size_t) { return B200CNN_COMM_FLAG_BYTES; }
extern "C" size_t b200cnn_comm_grads_offset(size_t flat_floats, int parity) {
    const size_t buf = align_up(flat_floats * sizeof(float), 4096);
    return B200CNN_COMM_FLAG_BYTES + buf * size_t(1 + (parity ? 1 : 0));
}
static size_t ll_offset(size_t flat_floats) { return B200CNN_COMM_FLAG_BYTES + 3 * align_up(flat_floats * sizeof(float), 4096); }
extern "C" size_t b200cnn_comm_region_bytes(size_t flat_floats) { return ll_offset(flat_floats) + ll::kTotalBytes; }

extern "C" b200cnn_comm* b200cnn_comm_create(int rank, int world, void* const* region_ptrs, void* multicast_ptr,
                                             size_t flat_floats) {
    if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world || !region_ptrs || flat_floats == 0 || (flat_floats & 3)) {
        set_error(kInvalidValue, "comm_create: need 1 <= world <= %d, 0 <= rank < world, flat_floats %% 4 == 0", kMaxRanks);
        return nullptr;
    }
    auto* c = new (std::nothrow) b200cnn_comm;
    if (!c) return nullptr;
    c->flat_floats = flat_floats;
    c->view.rank = rank;
    c->view.world = world;
    c->view.emulate = 0;
    for (int r = 0; r < kMaxRanks; r++) c->view.region[r] = static_cast<char*>(region_ptrs[r < world ? r : rank]);
    c->view.mc_region = static_cast<char*>(multicast_ptr);
    c->view.params_off = b200cnn_comm_params_offset(flat_floats);
    c->view.grads_off[0] = b200cnn_comm_grads_offset(flat_floats, 0);
    c->view.grads_off[1] = b200cnn_comm_grads_offset(flat_floats, 1);
    c->view.ll_off = ll_offset(flat_floats);
    const size_t local_bytes = (kMaxBlocks + 16) * sizeof(uint32_t) + ll::kChannels * sizeof(ll::Counters);
    if (cudaMalloc(&c->local, local_bytes) != cudaSuccess || cudaMemset(c->local, 0, local_bytes) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_wait, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) != cudaSuccess) {
        set_error(int(cudaGetLastError()), "comm_create: CUDA resource creation failed");
        delete c;
        return nullptr;
    }
    c->view.seq = c->local;
    c->view.errflag = c->local + kMaxBlocks;
    c->view.ll_ctr = reinterpret_cast<ll::Counters*>(c->local + kMaxBlocks + 16);
    return c;
}

extern "C" void b200cnn_comm_destroy(b200cnn_comm* c) {
    if (!c) return;
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->ev_wait) cudaEventDestroy(c->ev_wait);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->local) cudaFree(c->local);
    delete c;
}

extern "C" int b200cnn_comm_rank(const b200cnn_comm* c) { return c->view.rank; }
extern "C" int b200cnn_comm_world(const b200cnn_comm* c) { return c->view.world; }
extern "C" int b200cnn_comm_has_multicast(const b200cnn_comm* c) { return c->view.mc_region != nullptr; }
extern "C" float* b200cnn_comm_params(b200cnn_comm* c) {
    return reinterpret_cast<float*>(c->view.region[c->view.rank] + c->view.params_off);
}
extern "C" float* b200cnn_comm_grads(b200cnn_comm* c, int parity) {
    return reinterpret_cast<float*>(c->view.region[c->view.rank] + c->view.grads_off[parity ? 1 : 0]);
}
extern "C" int b200cnn_comm_set_algo(b200cnn_comm* c, int algo) {
    if (algo < B200CNN_ALGO_AUTO || algo > B200CNN_ALGO_LL_TWO_SHOT) return set_error(kInvalidValue, "comm_set_algo: unknown algorithm %d", algo);
    if (algo == B200CNN_ALGO_NVLS && !c->view.mc_region) return set_error(kInvalidValue, "comm_set_algo: no multicast mapping on this communicator");
    c->algo = algo;
    return 0;
}
extern "C" int b200cnn_comm_error(b200cnn_comm* c) {
    uint32_t flag = 0;
    if (cudaMemcpy(&flag, c->view.errflag, sizeof(flag), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return int(flag);
}

extern "C" int b200cnn_comm_allreduce_sgd(b200cnn_comm* c, size_t offset, size_t count, float lr, int algo, int parity,
                                          int max_blocks, b200cnn_stream stream) {
    return enqueue(c, offset, count, nullptr, lr, algo, parity, kOpSgd, 0, max_blocks, as_stream(stream));
}
extern "C" int b200cnn_comm_allreduce(b200cnn_comm* c, size_t offset, size_t count, int algo, int parity,
                                      int max_blocks, b200cnn_stream stream) {
    return enqueue(c, offset, count, nullptr, 0.0f, algo, parity, kOpReduceOnly, 0, max_blocks, as_stream(stream));
}
extern "C" int b200cnn_comm_broadcast_params(b200cnn_comm* c, int root, b200cnn_stream stream) {
    if (root < 0 || root >= c->view.world) return set_error(kInvalidValue, "comm_broadcast_params: bad root %d", root);
    return enqueue(c, 0, c->flat_floats, nullptr, 0.0f, B200CNN_ALGO_ONE_SHOT, 0, kOpBroadcast, root, 32, as_stream(stream));
}
extern "C" int b200cnn_comm_barrier(b200cnn_comm* c, b200cnn_stream stream) {
    return launch(c->view, 1, 0, 0, nullptr, 0.0f, B200CNN_ALGO_ONE_SHOT, 0, kOpBarrier, 0, 1, as_stream(stream));
}

extern "C" int b200cnn_comm_emulate_allreduce_sgd_slot(int world, void* const* region_ptrs, size_t flat_floats, size_t offset,
                                                       size_t count, float lr, int algo, int parity, int blocks,
                                                       int slot_base, uint32_t* seq_scratch, b200cnn_stream stream) {
    if (world < 1 || world > kMaxRanks || (offset | count) & 3 || offset + count > flat_floats)
        return set_error(kInvalidValue, "comm_emulate: bad arguments");
    if (algo != B200CNN_ALGO_ONE_SHOT && algo != B200CNN_ALGO_TWO_SHOT)
        return set_error(kInvalidValue, "comm_emulate: only one-shot and two-shot can run without a switch");
    if (blocks < 1 || blocks * world > 148) return set_error(kInvalidValue, "comm_emulate: blocks*world must fit one wave (<= 148)");
    CommView v{};
    v.world = world;
    v.emulate = 1;
    for (int r = 0; r < kMaxRanks; r++) v.region[r] = static_cast<char*>(region_ptrs[r < world ? r : 0]);
    v.params_off = b200cnn_comm_params_offset(flat_floats);
    v.grads_off[0] = b200cnn_comm_grads_offset(flat_floats, 0);
    v.grads_off[1] = b200cnn_comm_grads_offset(flat_floats, 1);
    v.seq = seq_scratch;                               // [world][kMaxBlocks]
    v.errflag = seq_scratch + size_t(world) * kMaxBlocks;  // + 1 word
    return launch(v, world, offset / 4, count / 4, nullptr, lr, algo, parity, kOpSgd, 0, blocks, as_stream(stream), slot_base);
}

extern "C" int b200cnn_comm_emulate_allreduce_sgd(int world, void* const* region_ptrs, size_t flat_floats, size_t offset,
                                                  size_t count, float lr, int algo, int parity, int blocks,
                                                  uint32_t* seq_scratch, b200cnn_stream stream) {
    return b200cnn_comm_emulate_allreduce_sgd_slot(world, region_ptrs, flat_floats, offset, count, lr, algo, parity, blocks, 0,
                                                   seq_scratch, stream);
}

extern "C" int b200cnn_comm_wait_weights(b200cnn_comm* c, b200cnn_stream stream) {
    if (c->view.world == 1) return 0;
    for (int ch = 0; ch < 2; ch++)
        if (int rc = launch_ll_wait(c->view, 1, ch, as_stream(stream))) return rc;
    return 0;
}

extern "C" size_t b200cnn_comm_emulate_ll_scratch_bytes(int world) {
    return size_t(world) * ll::kChannels * sizeof(ll::Counters) + 16;
}

extern "C" int b200cnn_comm_emulate_ll(int world, void* const* region_ptrs, size_t flat_floats, size_t offset, size_t count,
                                       float lr, int two_shot, int parity, int blocks, int channel, int update, void* scratch,
                                       b200cnn_stream stream) {
    if (world < 2 || world > kMaxRanks || (offset | count) & 3 || offset + count > flat_floats || channel < 0 || channel > 1)
        return set_error(kInvalidValue, "comm_emulate_ll: bad arguments");
    if (blocks < 1 || blocks > kLLMaxBlocks || blocks * world > 148) return set_error(kInvalidValue, "comm_emulate_ll: blocks*world must fit one wave (<= 148)");
    const size_t per_src = two_shot ? ((count / 4 + world - 1) / world * 16 + 16) : count * sizeof(float);
    if (per_src > ll::slot_bytes(channel) / 2) return set_error(kInvalidValue, "comm_emulate_ll: slice does not fit the inbox of channel %d", channel);
    CommView v{};
    v.world = world;
    v.emulate = 1;
    for (int r = 0; r < kMaxRanks; r++) v.region[r] = static_cast<char*>(region_ptrs[r < world ? r : 0]);
    v.params_off = b200cnn_comm_params_offset(flat_floats);
    v.grads_off[0] = b200cnn_comm_grads_offset(flat_floats, 0);
    v.grads_off[1] = b200cnn_comm_grads_offset(flat_floats, 1);
    v.ll_off = ll_offset(flat_floats);
    v.ll_ctr = static_cast<ll::Counters*>(scratch);
    v.errflag = reinterpret_cast<uint32_t*>(static_cast<char*>(scratch) + size_t(world) * ll::kChannels * sizeof(ll::Counters));
    if (int rc = launch_ll(v, world, offset / 4, count / 4, nullptr, lr, two_shot, parity, update ? kOpSgd : kOpReduceOnly, channel, blocks, as_stream(stream))) return rc;
    return two_shot ? launch_ll_wait(v, world, channel, as_stream(stream)) : 0;
}

// ---- engine hooks ---------------------------------------------------------------------------
namespace b200cnn {

int comm_allreduce_sgd_slice(b200cnn_comm* c, size_t offset, size_t count, const float* lr_dev, cudaStream_t after) {
    B200CNN_CUDA(cudaEventRecord(c->ev_fork, after));
    B200CNN_CUDA(cudaStreamWaitEvent(c->stream, c->ev_fork, 0));
    c->forked = true;
    // Few blocks: the kernel shares the GPU with the rest of backward.  (B200CNN_COMM_BLOCKS overrides, <= 64.)
    static const int max_blocks = [] {
        const char* s = std::getenv("B200CNN_COMM_BLOCKS");
        const int v = s ? std::atoi(s) : 32;
        return v < 1 ? 1 : v > kMaxBlocks / 2 ? kMaxBlocks / 2 : v;
    }();
    return enqueue(c, offset, count, lr_dev, 0.0f, B200CNN_ALGO_AUTO, c->parity, kOpSgd, 0, max_blocks, c->stream, 0, 4, /*ch=*/0,
                   /*defer_wait=*/true);
}

// Start of a step: the previous step's two-shot push calls stored their shards of the new weights into the peers and
// left; this enqueues (on the communicator's stream, next to whatever `st` does first) the kernel that waits until
// all peers' shards have landed HERE.  comm_weights_ready() then makes `st` wait for it — call it right before the
// first kernel that reads parameters a bucket may have updated.
int comm_begin_step(b200cnn_comm* c, cudaStream_t st) {
    if (c->view.world == 1) return 0;
    B200CNN_CUDA(cudaEventRecord(c->ev_fork, st));
    B200CNN_CUDA(cudaStreamWaitEvent(c->stream, c->ev_fork, 0));
    for (int ch = 0; ch < 2; ch++)
        if (int rc = launch_ll_wait(c->view, 1, ch, c->stream)) return rc;
    B200CNN_CUDA(cudaEventRecord(c->ev_wait, c->stream));
    c->wait_pending = true;
    c->forked = true;
    return 0;
}

bool comm_update_is_deferred(const b200cnn_comm* c, size_t count, int ch) {
    return c->view.world > 1 && resolve_algo(c, B200CNN_ALGO_AUTO, count, kOpSgd, ch) == B200CNN_ALGO_LL_TWO_SHOT;
}

int comm_weights_ready(b200cnn_comm* c, cudaStream_t st) {
    if (!c->wait_pending) return 0;
    B200CNN_CUDA(cudaStreamWaitEvent(st, c->ev_wait, 0));
    c->wait_pending = false;
    return 0;
}

// Same operation enqueued directly on `st` (no stream hop), on the second flag channel so it may overlap a
// kernel that comm_allreduce_sgd_slice put on the communicator's own stream.  Used for the last bucket of a step,
// which sits on the critical path.
int comm_allreduce_sgd_inline(b200cnn_comm* c, size_t offset, size_t count, const float* lr_dev, cudaStream_t st) {
    // A small slice here is pure latency: one flag exchange (one-shot) with one float4 per thread beats the two
    // exchanges of two-shot / NVLS at every world size.  B200CNN_COMM_TAIL_ALGO (1 one-shot, 2 two-shot, 3 NVLS)
    // overrides the choice for experiments.
    static const int forced = [] {
        const char* s = std::getenv("B200CNN_COMM_TAIL_ALGO");
        return s ? std::atoi(s) : 0;
    }();
    int algo = forced > 0 ? forced : B200CNN_ALGO_AUTO;
    if (forced <= 0 && !ll_enabled() && count * sizeof(float) <= (32u << 10)) algo = B200CNN_ALGO_ONE_SHOT;
    return enqueue(c, offset, count, lr_dev, 0.0f, algo, c->parity, kOpSgd, 0, 16, st, kMaxBlocks / 2, /*per_thread=*/1, /*ch=*/1,
                   /*defer_wait=*/true);
}

int comm_join(b200cnn_comm* c, cudaStream_t st) {
    if (!c->forked) return 0;
    B200CNN_CUDA(cudaEventRecord(c->ev_join, c->stream));
    B200CNN_CUDA(cudaStreamWaitEvent(st, c->ev_join, 0));
    c->forked = false;
    return 0;
}

// The last flag slot is reserved for exchanges done inside other kernels (phase 0 only, one-shot, no exit barrier:
// gradients are double-buffered by step parity).
void comm_tail_xchg(const b200cnn_comm* c, TailXchg* out) {
    out->chan.world = c->view.world;
    out->chan.rank = c->view.rank;
    out->chan.ch = 2;
    for (int r = 0; r < kMaxRanks; r++) out->chan.region[r] = c->view.region[r];
    out->chan.inbox_off = c->view.ll_off;
    out->chan.ctr = c->view.ll_ctr + 2;
    out->chan.errflag = c->view.errflag;
    out->params_off = c->view.params_off;
}

float* comm_current_grads(b200cnn_comm* c) { return b200cnn_comm_grads(c, c->parity); }
void comm_flip_parity(b200cnn_comm* c) { c->parity ^= 1; }
int comm_parity(const b200cnn_comm* c) { return c->parity; }

}  // namespace b200cnn
