// Training engine: owns the device-side state of one network replica and runs whole mini-batches.
//
// It replaces the per-sample driver of the reference — Layer_setInputs, Layer_learnOutputs,
// Layer_getErrorTotal and Layer_update called from main() once per image
// (/root/reference/cnn.c:451-474, same loop in CUDAcnn.cu:451-480) — by a fixed launch sequence over
// a batch, optionally captured as one CUDA graph so a step costs one launch on the host.
//
// Memory: all parameters live in ONE flat fp32 buffer and all gradients in another, each tensor
// starting on a 128-byte boundary.  Single GPU: the update is fused into backward (the upper layers on the
// side stream as soon as nothing reads their weights any more, the first layer inside its own weight-gradient
// kernel).  Data parallel: the buffers live in the communicator's NVLink-mapped region and the fused
// all-reduce + SGD runs per BUCKET of adjacent layer slices (>= 512 KiB, or whatever is left at the first
// layer), each launched as soon as its last gradient exists.
//
// Topologies with their own kernel chains (detected in b200cnn_engine_create, reported by b200cnn_engine_conv_path):
//   conv(1->16, s2) -> conv(16->32, s2) -> FC...     both forward convolutions in ONE root kernel (tc_conv12.cu), tcgen05
//                                                    implicit GEMMs for conv2 forward / backward-data / weight gradient,
//                                                    tcgen05 weight gradient for conv1 with the update (or the exchange
//                                                    with the peers) in its last block;
//   conv -> pool -> conv -> pool -> FC...            fused conv1 + pool, one-load halo convolutions (tc_conv_halo.cu),
//                                                    stride-1 tcgen05 weight gradient, conv1's weight gradient from the
//                                                    pooled gradient;
//   anything else                                    the generic CHW kernels of conv.cu / pool.cu and the tcgen05 / SIMT GEMMs.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "backend.cuh"
#include "common.cuh"
#include "../comm/comm.cuh"
#include "b200cnn_comm.h"

namespace b200cnn {

constexpr size_t kAlignFloats = 32;  // 128 B
static inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct EngineLayer {
    b200cnn_layer_desc d{};
    int nnodes = 0;
    int act = B200CNN_ACT_NONE;
    long w_off = -1, b_off = -1;
    long nw = 0, nb = 0;
    float* a = nullptr;     // activations  [max_batch][ld]
    float* dz = nullptr;    // d loss / d pre-activation [max_batch][ld]
    int32_t* argmax = nullptr;
    int ld = 0;             // row stride of a / dz (nnodes, except the padded softmax layer)
};

}  // namespace b200cnn

using namespace b200cnn;

struct b200cnn_engine {
    std::vector<EngineLayer> layers;
    int max_batch = 0;
    size_t dense_params = 0, flat_floats = 0;
    float* params = nullptr;
    float* grads = nullptr;
    bool own_buffers = false;
    void* arena = nullptr;       // activations, dz, scratch, metrics, hyper-parameters
    float* scratch = nullptr;   // partial sums of the conv weight-gradient kernels (layers >= 2)
    float* scratch1 = nullptr;  // same for layer 1, which may run concurrently with layer 2's
    size_t scratch_bytes = 0;
    float* metrics = nullptr;    // [4]
    float* lr_dev = nullptr;     // [1], read by the graph-captured update
    float lr_host_cached = -1.0f;
    // last forward
    const void* last_images = nullptr;
    int last_is_u8 = 0;
    int last_batch = 0;
    // Graph cache: a few entries, so a caller can rotate between staging buffers (copy batch i+1 while batch i
    // trains).  Key = (buffers, batch, input type, stream, backend); one executable per gradient-buffer parity.
    struct GraphEntry {
        cudaGraphExec_t exec[2] = {nullptr, nullptr};
        const void* images = nullptr;
        const uint8_t* labels = nullptr;
        int batch = 0, is_u8 = 0, backend = -1, update = 1;
        cudaStream_t stream = nullptr;
        unsigned long last_use = 0;
    };
    static constexpr int kGraphEntries = 4;
    GraphEntry graphs[kGraphEntries];
    unsigned long graph_clock = 0;
    b200cnn_comm* comm = nullptr;
    int launches_per_step = 0;
    cudaStream_t side = nullptr;          // weight-gradient kernels run here, next to the data-gradient chain
    cudaEvent_t ev_dz = nullptr, ev_data = nullptr, ev_side = nullptr;
    // Data-parallel update buckets: consecutive layers' slices are merged until a bucket holds kBucketFloats,
    // so the number of cross-GPU flag exchanges per step is 2 instead of one per layer.
    long bucket_lo = -1, bucket_hi = -1;
    long tail_lo = -1, tail_hi = -1;  // last bucket of the step: launched on the main stream after the join
    // host-fed steps (b200cnn_engine_train_step_host): two device staging slots + a copy stream, so batch i+1 crosses
    // PCIe while batch i trains
    uint8_t* stage_x[2] = {nullptr, nullptr};
    uint8_t* stage_y[2] = {nullptr, nullptr};
    size_t stage_item = 0;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_consumed[2] = {nullptr, nullptr};
    int host_slot = 0;
    int wait_layer = 1;       // data parallel: first layer whose parameters a deferred (two-shot push) update may write
    bool last_fast = false;   // layout of layer 1's activations in the last forward (NHWC when true)
    bool fast_chain = false;  // conv(C_in=1) -> tc conv3x3 -> FC pattern: NHWC activations + tcgen05 convolutions
    bool pool_chain = false;  // conv(C_in=1) -> pool -> tc conv3x3 (stride 1) -> pool -> FC: the same, all NHWC up to the last pool
    unsigned nhwc_mask = 0;   // bit li: layer li's activations of the last forward are NHWC on the device
    bool pool_fused = false;  // pool chain: conv1 + ReLU + pool as one kernel, conv1's weight gradient from the pooled gradient
    bool keep_full = true;    // ... and whether that kernel also stores the un-pooled activations (forward(): yes, steps: no)
    bool l1_stale = false;    // the last forward did not store layer 1's activations
};

namespace {

// Gradient buffer of the step being enqueued: the communicator's double buffer when attached.
float* cur_grads(b200cnn_engine* e) { return e->comm ? comm_current_grads(e->comm) : e->grads; }

b200cnn_conv_shape conv_shape(const EngineLayer& prev, const EngineLayer& l, int batch) {
    b200cnn_conv_shape s;
    s.n = batch;
    s.c = prev.d.depth;
    s.h = prev.d.height;
    s.w = prev.d.width;
    s.oc = l.d.depth;
    s.oh = l.d.height;
    s.ow = l.d.width;
    s.k = l.d.k;
    s.pad = l.d.pad;
    s.stride = l.d.stride;
    return s;
}

// The tensor-core convolution chain is used when the topology allows it and the backend is not pinned
// to the CUDA-core kernels.
bool use_fast_chain(const b200cnn_engine* e) { return e->fast_chain && current_backend() != B200CNN_BACKEND_SIMT; }
bool use_pool_chain(const b200cnn_engine* e) { return e->pool_chain && current_backend() != B200CNN_BACKEND_SIMT; }

// True when the last three layers are FC(tanh) -> FC(tanh) -> FC(softmax) of sizes the fused head kernel takes.
bool use_head(const b200cnn_engine* e, int batch) {
    // Measured on B200 at batch 256: the fused kernel (2 CTAs, four serial GEMM phases, all epilogue traffic on two
    // SMs) takes 45 us against 26 us for the five separate kernels it replaces (13-26 CTAs each), so it is opt-in.
    static const bool enabled = [] {
        const char* v = getenv("B200CNN_HEAD");
        return v && v[0] == '1';
    }();
    if (!enabled) return false;
    const auto& L = e->layers;
    const size_t n = L.size();
    if (n < 4 || current_backend() == B200CNN_BACKEND_SIMT) return false;
    for (size_t i = n - 3; i < n; i++)
        if (L[i].d.type != B200CNN_LAYER_FULL) return false;
    return tc_head_supported(batch, L[n - 3].nnodes, L[n - 2].nnodes, L[n - 1].nnodes, L[n - 1].ld);
}

// True when the last three layers are FC(tanh) -> FC(tanh) -> FC(softmax) on top of a float layer and the cluster
// kernel of tc_chain.cu takes the shape: forward of the three layers, loss gradient and the three data gradients then
// run as ONE launch.
bool use_chain(const b200cnn_engine* e, int batch) {
    const auto& L = e->layers;
    const size_t n = L.size();
    if (n < 5 || !tc_fc_chain_enabled()) return false;
    for (size_t i = n - 3; i < n; i++)
        if (L[i].d.type != B200CNN_LAYER_FULL) return false;
    if (L[n - 4].d.type == B200CNN_LAYER_INPUT || L[n - 4].ld != L[n - 4].nnodes) return false;
    return tc_fc_chain_supported(batch, L[n - 4].nnodes, L[n - 3].nnodes, L[n - 2].nnodes, L[n - 1].nnodes, L[n - 1].ld);
}

// Same topology test for the cluster kernel that fuses only the middle (fc -> softmax/loss -> two data-gradient GEMMs).
bool use_mid(const b200cnn_engine* e, int batch) {
    const auto& L = e->layers;
    const size_t n = L.size();
    if (n < 4 || !tc_fc_mid_enabled()) return false;
    for (size_t i = n - 3; i < n; i++)
        if (L[i].d.type != B200CNN_LAYER_FULL) return false;
    return tc_fc_mid_supported(batch, L[n - 3].nnodes, L[n - 2].nnodes, L[n - 1].nnodes, L[n - 1].ld);
}

// The tensor-core convolutions gather their weights into shared memory before griddepcontrol.wait: inside a step the kernel
// in front of them on the stream is always another layer's kernel (never an update), and that kernel released its
// dependents only after its own wait, i.e. after everything older — including the previous step's update — was complete.
// (A data-parallel update arriving from the peers is ordered by an event, a full dependency: the kernel does not start early.)
constexpr int kEarlyW = 1;

// Forward through layers [1, stop).
int enqueue_forward(b200cnn_engine* e, const void* images, int is_u8, const uint8_t* labels, int batch,
                    cudaStream_t st, int* launches, size_t stop = ~size_t(0)) {
    auto& L = e->layers;
    const bool fast = use_fast_chain(e);
    const bool poolc = use_pool_chain(e);
    e->last_fast = fast;
    e->nhwc_mask = fast ? 2u : poolc ? 14u : 0u;  // NHWC layers: 1, or 1-3 (the second pool writes CHW rows for the FC layer)
    const bool fusedp = poolc && e->pool_fused;
    e->l1_stale = fusedp && !e->keep_full;
    // Reference topology: both convolutions in ONE kernel, the root of the captured step (tc_conv12.cu)
    bool fused12 = false;
    if (fast && is_u8 && stop > 2 && L.size() > 3 && tc_conv12_fwd_enabled() && L[1].d.type == B200CNN_LAYER_CONV &&
        L[2].d.type == B200CNN_LAYER_CONV) {
        const b200cnn_conv_shape s1 = conv_shape(L[0], L[1], batch), s2 = conv_shape(L[1], L[2], batch);
        fused12 = tc_conv12_fwd_supported(&s1, &s2);
    }
    for (size_t li = 1; li < L.size() && li < stop; li++) {
        EngineLayer& l = L[li];
        const EngineLayer& p = L[li - 1];
        const void* in = (li == 1) ? images : static_cast<const void*>(p.a);
        const int in_u8 = (li == 1) ? is_u8 : 0;
        int rc = 0;
        // Data parallel: from this layer on the parameters may have been written by the peers (two-shot push of the
        // previous step); the wait for their arrival was enqueued on the communicator's stream at the start of the step.
        if (e->comm && int(li) >= e->wait_layer && l.nw > 0)
            if ((rc = comm_weights_ready(e->comm, st))) return rc;
        switch (l.d.type) {
            case B200CNN_LAYER_CONV: {
                const b200cnn_conv_shape s = conv_shape(p, l, batch);
                if (fused12 && li == 2) continue;  // computed by the kernel of layer 1
                if (fused12 && li == 1) {
                    EngineLayer& l2 = L[2];
                    const b200cnn_conv_shape s2 = conv_shape(l, l2, batch);
                    if (e->comm && 2 >= e->wait_layer)  // the kernel reads layer 2's parameters as well
                        if ((rc = comm_weights_ready(e->comm, st))) return rc;
                    rc = tc_conv12_fwd(static_cast<const uint8_t*>(in), e->params + l.w_off, e->params + l.b_off, l.a, l.act,
                                       e->params + l2.w_off, e->params + l2.b_off, l2.a, l2.ld, l2.act, &s, &s2, st);
                } else if (fusedp && li == 1)  // convolution + ReLU + the pool above it in one kernel (layer 2 is skipped below)
                    rc = conv1_pool_fwd(in, in_u8, e->params + l.w_off, e->params + l.b_off, e->keep_full ? l.a : nullptr, L[2].a,
                                        L[2].argmax, &s, l.act, st);
                else if ((fast || poolc) && li == 1)  // first conv: CUDA cores (1 input channel), NHWC output for the TMA loads below
                    rc = simt_conv2d_fwd(in, in_u8, e->params + l.w_off, e->params + l.b_off, l.a, &s, l.act, 1, st);
                else if (poolc && li == 3)  // 196 pixels per image: two bands of the implicit GEMM, NHWC output for the pool
                    rc = tc_conv3x3_fwd(static_cast<const float*>(in), e->params + l.w_off, e->params + l.b_off, l.a, 0, &s, l.act, st, 1,
                                        kEarlyW);
                else if (fast && li == 2)  // implicit GEMM on tcgen05, output CHW-flattened for the FC layer
                    rc = tc_conv3x3_fwd(static_cast<const float*>(in), e->params + l.w_off, e->params + l.b_off, l.a, l.ld, &s,
                                        l.act, st, 0, kEarlyW);
                else
                    rc = b200cnn_conv2d_fwd(in, in_u8, e->params + l.w_off, e->params + l.b_off, l.a, &s, l.act, st);
                break;
            }
            case B200CNN_LAYER_POOL:
                if (in_u8) return set_error(kInvalidValue, "engine: a pool layer cannot read uint8 input directly");
                if (fusedp && li == 2) continue;  // computed by the first convolution's kernel
                if (poolc)  // NHWC in; NHWC out, except the last pool, whose output an FC layer reads as CHW rows
                    rc = maxpool_nhwc_fwd(static_cast<const float*>(in), l.a, l.argmax, batch, p.d.depth, p.d.height, p.d.width, l.d.k,
                                          l.d.stride, li == 4 ? 1 : 0, st);
                else
                rc = b200cnn_maxpool_fwd(static_cast<const float*>(in), l.a, l.argmax, batch, p.d.depth, p.d.height,
                                         p.d.width, l.d.k, l.d.stride, st);
                break;
            case B200CNN_LAYER_FULL:
                if (in_u8) return set_error(kInvalidValue, "engine: a full layer cannot read uint8 input directly");
                if (li + 1 == L.size()) {  // output layer: GEMM + softmax + loss gradient + metrics
                    rc = b200cnn_linear_softmax_xent(static_cast<const float*>(in), e->params + l.w_off, e->params + l.b_off,
                                                     labels, l.a, labels ? l.dz : nullptr, labels ? e->metrics : nullptr,
                                                     batch, l.nnodes, p.nnodes, l.ld, st);
                    if (!rc && launches)
                        (*launches) += (current_backend() != B200CNN_BACKEND_SIMT && tc_linear_softmax_supported(batch, l.nnodes, p.nnodes, l.ld)) ? 0 : 1;
                    break;
                }
                rc = b200cnn_linear_fwd(static_cast<const float*>(in), e->params + l.w_off, e->params + l.b_off, l.a,
                                        batch, l.nnodes, p.nnodes, l.ld, l.act, st);
                break;
            default: return set_error(kInvalidValue, "engine: bad layer type %d", l.d.type);
        }
        if (rc) return rc;
        if (launches) (*launches)++;
    }
    return 0;
}

// Hook after a layer's gradients are complete (and its data-gradient kernel has been issued): the
// data-parallel path adds the layer's slice to the current bucket and launches the fused all-reduce + SGD
// for the bucket once it is big enough (or at the last layer), so it overlaps the rest of backward.  Layers
// are visited last-to-first and their slices are adjacent in the flat buffer, so a bucket is one range.
constexpr long kBucketFloats = 128 * 1024;  // 512 KiB

int layer_grads_ready(b200cnn_engine* e, const EngineLayer& l, bool last, cudaStream_t st, int* launches) {
    if (!e->comm) return 0;
    const long begin = l.w_off;
    const long end = l.b_off + long(round_up(size_t(l.nb), kAlignFloats));
    if (e->bucket_lo < 0) {
        e->bucket_lo = begin;
        e->bucket_hi = end;
    } else {
        e->bucket_lo = std::min(e->bucket_lo, begin);
        e->bucket_hi = std::max(e->bucket_hi, end);
    }
    if (!last && e->bucket_hi - e->bucket_lo < kBucketFloats) return 0;
    if (last) {  // deferred: runs on the main stream right after the side stream has been joined
        e->tail_lo = e->bucket_lo;
        e->tail_hi = e->bucket_hi;
        e->bucket_lo = e->bucket_hi = -1;
        return 0;
    }
    const int rc = comm_allreduce_sgd_slice(e->comm, size_t(e->bucket_lo), size_t(e->bucket_hi - e->bucket_lo), e->lr_dev, st);
    e->bucket_lo = e->bucket_hi = -1;
    if (rc) return rc;
    if (launches) (*launches)++;
    return 0;
}

// Backward.  The data-gradient chain (dz of layer l -> dz of layer l-1) is the critical path and stays on
// `st`; every weight-gradient kernel only needs its own layer's dz, so it runs on the engine's side
// stream, concurrently with the rest of the chain.  (Inside a captured step these become parallel
// branches of the CUDA graph.)  Order on the side stream per layer:
//     wait(dz ready) -> dW/db kernel(s) -> wait(this layer's data gradient issued) -> fused update hook
// The hook must come after the data-gradient kernel because a two-shot / NVLS update may overwrite the
// layer's weights on this GPU from a peer.  The step's LAST bucket is the exception: it is launched on `st`
// itself after the side stream has been joined (second flag channel of the communicator), because it is the tail
// of the step and every stream hop in front of it is exposed latency.
int enqueue_sgd_range(b200cnn_engine* e, size_t begin, size_t end, cudaStream_t st, int* launches);

static bool fused_tail_requested() {
    static const bool on = [] {
        const char* v = getenv("B200CNN_FUSED_TAIL");  // B200CNN_FUSED_TAIL=0: separate collective kernel after backward
        return !(v && v[0] == '0');
    }();
    return on;
}

// dx_done: the data gradients INTO the last `dx_done` layers' inputs already exist (fused head / chain kernel).
int enqueue_backward(b200cnn_engine* e, const void* images, int is_u8, int batch, cudaStream_t st, int* launches,
                     int dx_done = 0, bool fuse_update = false) {
    auto& L = e->layers;
    float* grads = cur_grads(e);
    const bool fast = use_fast_chain(e);
    const bool poolc = use_pool_chain(e);
    const bool fusedp = poolc && e->pool_fused;
    const size_t second = poolc ? 3 : 2;  // the first parameterised layer above layer 1
    cudaStream_t side = e->side;
    e->bucket_lo = e->bucket_hi = -1;
    e->tail_lo = e->tail_hi = -1;
    // Data parallel, tensor-core chain: the first convolution's weight-gradient kernel all-reduces its 160 numbers
    // itself in its last block (push protocol, csrc/comm/xchg_device.cuh) and the second convolution's slice goes
    // out on the side stream as soon as it exists, so no collective kernel is left on the main stream after the last
    // kernel of backward.
    TailXchg xchg;
    bool fused_tail = false;
    if (e->comm && fast && fused_tail_requested() && tc_conv1_bwd_weight_enabled() && is_u8 && L.size() > 3 &&
        L[1].d.type == B200CNN_LAYER_CONV && L[0].d.depth == 1 && L[2].d.type == B200CNN_LAYER_CONV) {
        const b200cnn_conv_shape s1 = conv_shape(L[0], L[1], batch);
        if (tc_conv1_bwd_weight_supported(&s1, nhwc_bwd_weight_chunks(batch, nullptr))) {
            comm_tail_xchg(e->comm, &xchg);
            fused_tail = xchg.world() > 1;
        }
    }
    // Reference topology: the second convolution's data-gradient kernel also produces the first convolution's weight
    // gradient (tc_conv_tail.cu) — dz1 never leaves the SM and the step loses its last kernel boundary.
    bool c1_in_dx = false;
    if (fast && is_u8 && tc_conv_tail_enabled() && L.size() > 3 && L[1].d.type == B200CNN_LAYER_CONV && L[2].d.type == B200CNN_LAYER_CONV &&
        (!e->comm || fused_tail)) {
        const b200cnn_conv_shape s1 = conv_shape(L[0], L[1], batch), s2 = conv_shape(L[1], L[2], batch);
        c1_in_dx = tc_conv_tail_supported(&s2, &s1, nhwc_bwd_weight_chunks(batch, nullptr)) && L[1].act == B200CNN_ACT_RELU;
    }
    for (size_t li = L.size() - 1; li >= 1; li--) {
        EngineLayer& l = L[li];
        EngineLayer& p = L[li - 1];
        const void* in = (li == 1) ? images : static_cast<const void*>(p.a);
        const int in_u8 = (li == 1) ? is_u8 : 0;
        int rc = 0;
        if (l.d.type == B200CNN_LAYER_POOL) {
            if (fusedp && li == 2) continue;  // the first convolution takes its weight gradient from THIS layer's gradient
            if (li > 1) {
                if (poolc)  // gradient arrives CHW (from the FC layer) at the last pool, NHWC elsewhere; leaves NHWC
                    rc = maxpool_nhwc_bwd(l.dz, li == 4 ? 1 : 0, l.argmax, p.a, p.act, p.dz, batch, p.d.depth, p.d.height, p.d.width,
                                          l.d.k, l.d.stride, st);
                else
                rc = b200cnn_maxpool_bwd(l.dz, l.argmax, p.a, p.act, p.dz, batch, p.d.depth, p.d.height, p.d.width, l.d.k,
                                         l.d.stride, st);
                if (rc) return rc;
                if (launches) (*launches)++;
            }
            continue;
        }
        if (l.d.type != B200CNN_LAYER_FULL && l.d.type != B200CNN_LAYER_CONV)
            return set_error(kInvalidValue, "engine: bad layer type %d", l.d.type);

        // l.dz is complete on `st` here: fork the weight-gradient work.  The first parameterised layer has no
        // data gradient to compute, so its weight gradient IS the tail of the critical path: it stays on `st`
        // instead of queueing behind the other weight-gradient kernels on the side stream.
        const bool tail = (li == 1);
        cudaStream_t ws = tail ? st : side;
        if (!tail) {
            B200CNN_CUDA(cudaEventRecord(e->ev_dz, st));
            B200CNN_CUDA(cudaStreamWaitEvent(side, e->ev_dz, 0));
        }
        if (l.d.type == B200CNN_LAYER_FULL) {
            rc = b200cnn_linear_bwd_weight(l.dz, static_cast<const float*>(in), grads + l.w_off, grads + l.b_off, 0, batch,
                                           l.nnodes, p.nnodes, l.ld, ws);
            if (launches) (*launches)++;
        } else {
            const b200cnn_conv_shape s = conv_shape(p, l, batch);
            const bool nhwc_path = (fast && li <= 2) || (poolc && (li == 1 || li == 3));
            const bool fuse = fuse_update && li == 1;  // single GPU: this kernel's last block also applies the update
            if (c1_in_dx && li == 1) {
                // already produced (and applied / exchanged) by the kernel that computed this layer's dz
            } else if (fusedp && li == 1) {
                rc = conv1_pool_bwd_weight(in, in_u8, L[2].dz, L[2].a, L[2].argmax, l.act, grads + l.w_off, grads + l.b_off, 0, e->scratch1,
                                           reinterpret_cast<unsigned*>(e->lr_dev + 8), &s, ws, fuse ? e->params + l.w_off : nullptr,
                                           fuse ? e->params + l.b_off : nullptr, e->lr_dev);
            } else if (nhwc_path) {
                const bool xc = fused_tail && li == 1;  // this kernel exchanges its result with the peers itself
                rc = nhwc_conv_bwd_weight(in, in_u8, l.dz, /*dz_chw=*/fast && li == 2, grads + l.w_off, grads + l.b_off, 0,
                                          li == 1 ? e->scratch1 : e->scratch, reinterpret_cast<unsigned*>(e->lr_dev + 8), &s, ws,
                                          fuse ? e->params + l.w_off : nullptr, fuse ? e->params + l.b_off : nullptr, e->lr_dev,
                                          xc ? &xchg : nullptr, long(l.w_off), long(l.b_off));
            }
            else if (b200cnn_conv2d_bwd_weight_scratch_bytes(&s) > e->scratch_bytes)
                rc = set_error(kInvalidValue, "engine: weight-gradient partial sums of layer %zu do not fit the scratch buffer", li);
            else
                rc = launch_conv2d_bwd_weight(l.dz, in, in_u8, grads + l.w_off, grads + l.b_off, 0,
                                              li == 1 ? e->scratch1 : e->scratch, &s, ws);
            // (the single-channel kernel reduces in its last block: one launch; every other path is kernel + reduce)
            if (launches) (*launches) += (c1_in_dx && li == 1) ? 0 : (nhwc_path && p.d.depth == 1) ? 1 : 2;
        }
        if (rc) return rc;

        // Data gradient for the layer below (critical path); the fused head kernel has already produced it for
        // the last two layers.
        if (li > 1 && li + size_t(dx_done) < L.size()) {
            if (l.d.type == B200CNN_LAYER_FULL) {
                rc = b200cnn_linear_bwd_data(l.dz, e->params + l.w_off, p.a, p.act, p.dz, batch, l.nnodes, p.nnodes, l.ld, st);
            } else {
                const b200cnn_conv_shape s = conv_shape(p, l, batch);
                if (fast && li == 2 && c1_in_dx) {
                    const EngineLayer& l1 = L[1];
                    const b200cnn_conv_shape s1 = conv_shape(L[0], l1, batch);
                    rc = tc_conv2dx_conv1dw(l.dz, /*dz_is_chw=*/1, e->params + l.w_off, p.a, p.act, /*dx=*/nullptr, &s,
                                            static_cast<const uint8_t*>(images), grads + l1.w_off, grads + l1.b_off, e->scratch1 + 32,
                                            reinterpret_cast<unsigned*>(e->lr_dev + 8), &s1, fuse_update ? e->params + l1.w_off : nullptr,
                                            fuse_update ? e->params + l1.b_off : nullptr, e->lr_dev, fused_tail ? &xchg : nullptr,
                                            long(l1.w_off), long(l1.b_off), kEarlyW, st);
                } else if (fast && li == 2)
                    rc = tc_conv3x3s2_bwd_data(l.dz, /*dz_is_chw=*/1, e->params + l.w_off, p.a, p.act, p.dz, &s, st, kEarlyW);
                else if (poolc && li == 3)  // the layer below is a pool: no activation derivative to apply
                    rc = tc_conv3x3s1_bwd_data(l.dz, e->params + l.w_off, nullptr, B200CNN_ACT_NONE, p.dz, &s, st, kEarlyW);
                else
                    rc = b200cnn_conv2d_bwd_data(l.dz, e->params + l.w_off, p.a, p.act, p.dz, &s, st);
            }
            if (rc) return rc;
            if (launches) (*launches)++;
        }
        if (fuse_update && li == second) {
            // Every layer above the first has its gradient (side stream) and nothing reads its weights any more once
            // this layer's data-gradient kernel (just issued on `st`) is done: update them now, next to the tail.
            B200CNN_CUDA(cudaEventRecord(e->ev_data, st));
            B200CNN_CUDA(cudaStreamWaitEvent(side, e->ev_data, 0));
            if ((rc = enqueue_sgd_range(e, size_t(l.w_off), e->flat_floats, side, launches))) return rc;
        }
        if (e->comm) {
            bool last = true;  // no parameterised layer below this one?
            for (size_t lj = 1; lj < li; lj++) last = last && L[lj].nw == 0;
            if (!last) {  // the update hook (on the side stream) follows everything issued on `st` for this layer
                B200CNN_CUDA(cudaEventRecord(e->ev_data, st));
                B200CNN_CUDA(cudaStreamWaitEvent(side, e->ev_data, 0));
            }
            if (fused_tail && li == 2) {
                // (anything the layers above left in the bucket goes first, in order, on the communicator's stream)
                if (e->bucket_lo >= 0) {
                    if ((rc = comm_allreduce_sgd_slice(e->comm, size_t(e->bucket_lo), size_t(e->bucket_hi - e->bucket_lo), e->lr_dev, side))) return rc;
                    e->bucket_lo = e->bucket_hi = -1;
                    if (launches) (*launches)++;
                }
                const long end = l.b_off + long(round_up(size_t(l.nb), kAlignFloats));
                if ((rc = comm_allreduce_sgd_inline(e->comm, size_t(l.w_off), size_t(end - l.w_off), e->lr_dev, side))) return rc;
                if (launches) (*launches)++;
            } else if (fused_tail && li == 1) {
                // exchanged inside its weight-gradient kernel
            } else if ((rc = layer_grads_ready(e, l, last, side, launches))) {
                return rc;
            }
        }
    }
    // Join: everything on the side stream is ordered before whatever follows on `st`.
    B200CNN_CUDA(cudaEventRecord(e->ev_side, side));
    B200CNN_CUDA(cudaStreamWaitEvent(st, e->ev_side, 0));
    if (e->comm && e->tail_lo >= 0) {
        // The step's last bucket is on the critical path: no stream hops, and it does not queue behind the big
        // bucket that may still be running on the communicator's stream (separate flag channel).
        const int rc = comm_allreduce_sgd_inline(e->comm, size_t(e->tail_lo), size_t(e->tail_hi - e->tail_lo), e->lr_dev, st);
        e->tail_lo = e->tail_hi = -1;
        if (rc) return rc;
        if (launches) (*launches)++;
    }
    return 0;
}

__global__ void sgd_dev_lr_kernel(float* __restrict__ w, const float* __restrict__ g, size_t n4,
                                  const float* __restrict__ lr_dev) {
    pdl_wait();
    pdl_launch_dependents();
    const float lr = *lr_dev;
    float4* w4 = reinterpret_cast<float4*>(w);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += size_t(gridDim.x) * blockDim.x) {
        float4 a = w4[i];
        const float4 d = g4[i];
        a.x -= lr * d.x;
        a.y -= lr * d.y;
        a.z -= lr * d.z;
        a.w -= lr * d.w;
        w4[i] = a;
    }
}

// Experiment (B200CNN_GRAPH_ROOT=1): an empty kernel in front of the step.  Measured: the second kernel node of a launched
// graph starts no earlier than ~6.2 us after the root started, whatever the root does (empty root: conv1 forward starts at
// 6.9 us; without it conv1 forward runs 0 - 4.5 us and conv2 forward starts at 6.1 us); step time is the same either way.
__global__ void graph_root_kernel() {
    pdl_wait();
    pdl_launch_dependents();
}
static bool graph_root_requested() {
    static const bool on = [] {
        const char* v = getenv("B200CNN_GRAPH_ROOT");
        return v && v[0] == '1';
    }();
    return on;
}

int enqueue_sgd_range(b200cnn_engine* e, size_t begin, size_t end, cudaStream_t st, int* launches) {
    const size_t n4 = (end - begin) / 4;
    const int blocks = int(std::min<size_t>((n4 + 255) / 256, 148 * 4));
    B200CNN_CUDA(launch_kernel(sgd_dev_lr_kernel, dim3(blocks), dim3(256), 0, st, e->params + begin,
                               static_cast<const float*>(cur_grads(e) + begin), n4, static_cast<const float*>(e->lr_dev)));
    if (launches) (*launches)++;
    return 0;
}

// True when backward can apply the update itself (single GPU, tensor-core conv chain): layers >= 2 on the side
// stream as soon as their weights are no longer read, layer 1 inside its weight-gradient kernel.
bool can_fuse_update(const b200cnn_engine* e) {
    return !e->comm && ((use_fast_chain(e) && e->layers[2].nw > 0) || (use_pool_chain(e) && e->layers[3].nw > 0));
}

int enqueue_update(b200cnn_engine* e, cudaStream_t st, int* launches) {
    if (e->comm) return comm_join(e->comm, st);  // per-layer fused updates already enqueued
    return enqueue_sgd_range(e, 0, e->flat_floats, st, launches);
}

int set_lr(b200cnn_engine* e, float lr, cudaStream_t st) {
    if (lr == e->lr_host_cached) return 0;
    // A 4-byte pageable copy is staged by the runtime, so `lr` may go out of scope immediately.
    B200CNN_CUDA(cudaMemcpyAsync(e->lr_dev, &lr, sizeof(float), cudaMemcpyHostToDevice, st));
    e->lr_host_cached = lr;
    return 0;
}

}  // namespace

extern "C" b200cnn_engine* b200cnn_engine_create(const b200cnn_layer_desc* layers, int nlayers, int max_batch) {
    if (!layers || nlayers < 2 || max_batch < 1 || layers[0].type != B200CNN_LAYER_INPUT ||
        layers[nlayers - 1].type != B200CNN_LAYER_FULL) {
        set_error(kInvalidValue, "engine_create: need input ... full layers and max_batch >= 1");
        return nullptr;
    }
    auto* e = new b200cnn_engine;
    e->max_batch = max_batch;
    size_t off = 0, arena_floats = 0, scratch_bytes = 0;
    for (int i = 0; i < nlayers; i++) {
        EngineLayer l;
        l.d = layers[i];
        l.nnodes = l.d.depth * l.d.width * l.d.height;
        l.ld = l.nnodes;
        if (i > 0) {
            const EngineLayer& p = e->layers.back();
            if (l.d.type == B200CNN_LAYER_CONV) {
                l.nw = long(l.d.depth) * p.d.depth * l.d.k * l.d.k;
                l.nb = l.d.depth;
                l.act = B200CNN_ACT_RELU;
                // The number of partial sums is NOT monotonic in the batch (ceil(n / ceil(n / 296)): 296 chunks at
                // n = 296, 149 at n = 297), so size the scratch for the worst legal batch, not for max_batch.
                b200cnn_conv_shape s = conv_shape(p, l, max_batch);
                for (int n = 1; n <= max_batch; n++) {
                    s.n = n;
                    scratch_bytes = std::max(scratch_bytes, b200cnn_conv2d_bwd_weight_scratch_bytes(&s));
                }
            } else if (l.d.type == B200CNN_LAYER_FULL) {
                l.nw = long(l.nnodes) * p.nnodes;
                l.nb = l.nnodes;
                l.act = (i == nlayers - 1) ? B200CNN_ACT_NONE : B200CNN_ACT_TANH;
                if (i == nlayers - 1) l.ld = int(round_up(size_t(l.nnodes), 16));  // 64-byte rows for the logits
            } else if (l.d.type == B200CNN_LAYER_POOL) {
                if (l.d.depth != p.d.depth) {
                    set_error(kInvalidValue, "engine_create: pool layer %d changes depth", i);
                    delete e;
                    return nullptr;
                }
            }
            if (l.nw) {
                l.w_off = long(off);
                off += round_up(size_t(l.nw), kAlignFloats);
                l.b_off = long(off);
                off += round_up(size_t(l.nb), kAlignFloats);
                e->dense_params += size_t(l.nw + l.nb);
            }
        }
        arena_floats += 2 * round_up(size_t(max_batch) * l.ld, kAlignFloats);
        if (l.d.type == B200CNN_LAYER_POOL) arena_floats += round_up(size_t(max_batch) * l.nnodes, kAlignFloats);
        e->layers.push_back(l);
    }
    e->flat_floats = off;
    if (nlayers >= 4 && e->layers[1].d.type == B200CNN_LAYER_CONV && e->layers[0].d.depth == 1 &&
        e->layers[2].d.type == B200CNN_LAYER_CONV && e->layers[3].d.type == B200CNN_LAYER_FULL) {
        const b200cnn_conv_shape s1 = conv_shape(e->layers[0], e->layers[1], max_batch);
        const b200cnn_conv_shape s2 = conv_shape(e->layers[1], e->layers[2], max_batch);
        e->fast_chain = tc_conv3x3_fwd_supported(&s2) && tc_conv3x3s2_bwd_data_supported(&s2) &&
                        nhwc_bwd_weight_supported(&s1, 1) && nhwc_bwd_weight_supported(&s2, 0) && s1.oc % 16 == 0 &&
                        tc_linear_bwd_data_supported(max_batch, e->layers[3].nnodes, e->layers[2].nnodes, e->layers[3].ld);
        if (e->fast_chain) {
            // worst case over every legal batch (same non-monotonic shape as above, period 148)
            b200cnn_conv_shape t1 = s1, t2 = s2;
            for (int n = 1; n <= max_batch; n++) {
                t1.n = t2.n = n;
                scratch_bytes = std::max(scratch_bytes, nhwc_bwd_weight_scratch_floats(&t1) * sizeof(float) + 128);
                scratch_bytes = std::max(scratch_bytes, nhwc_bwd_weight_scratch_floats(&t2) * sizeof(float) + 128);
            }
        }
    }
    // The pooled variant conv -> pool -> conv -> pool -> FC: first convolution on the CUDA cores (one input channel),
    // the second one (stride 1, 14x14) as a banded implicit GEMM on tcgen05, NHWC between them.
    if (nlayers >= 6 && e->layers[0].d.depth == 1 && e->layers[1].d.type == B200CNN_LAYER_CONV &&
        e->layers[2].d.type == B200CNN_LAYER_POOL && e->layers[3].d.type == B200CNN_LAYER_CONV &&
        e->layers[4].d.type == B200CNN_LAYER_POOL && e->layers[5].d.type == B200CNN_LAYER_FULL) {
        const b200cnn_conv_shape s1 = conv_shape(e->layers[0], e->layers[1], max_batch);
        const b200cnn_conv_shape s3 = conv_shape(e->layers[2], e->layers[3], max_batch);
        constexpr size_t kSmemLimit = 200 * 1024;
        e->pool_chain = s1.oc % 16 == 0 && tc_conv3x3_fwd_supported(&s3) && tc_conv3x3s1_bwd_data_supported(&s3) &&
                        nhwc_bwd_weight_supported(&s1, 1) && nhwc_bwd_weight_supported(&s3, 0) &&
                        nhwc_bwd_weight_smem_bytes(&s1) <= kSmemLimit && nhwc_bwd_weight_smem_bytes(&s3) <= kSmemLimit;
        if (e->pool_chain) {
            b200cnn_conv_shape t1 = s1, t3 = s3;
            for (int n = 1; n <= max_batch; n++) {
                t1.n = t3.n = n;
                scratch_bytes = std::max(scratch_bytes, nhwc_bwd_weight_scratch_floats(&t1) * sizeof(float) + 128);
                scratch_bytes = std::max(scratch_bytes, nhwc_bwd_weight_scratch_floats(&t3) * sizeof(float) + 128);
            }
            static const bool fuse_env = [] {
                const char* v = getenv("B200CNN_POOL_FUSED");  // B200CNN_POOL_FUSED=0: separate conv / pool kernels (A/B timing)
                return !(v && v[0] == '0');
            }();
            e->pool_fused = fuse_env && e->layers[1].act == B200CNN_ACT_RELU && conv1_pool_supported(&s1, e->layers[2].d.k, e->layers[2].d.stride);
        }
    }
    e->scratch_bytes = round_up(scratch_bytes, 512);
    const size_t arena_bytes = arena_floats * sizeof(float) + 2 * e->scratch_bytes + 1024;  // tail: metrics, lr, counters
    if (cudaMalloc(&e->arena, arena_bytes) != cudaSuccess || cudaMemset(e->arena, 0, arena_bytes) != cudaSuccess) {
        set_error(int(cudaGetLastError()), "engine_create: cannot allocate %zu bytes of activations", arena_bytes);
        delete e;
        return nullptr;
    }
    float* cur = static_cast<float*>(e->arena);
    for (EngineLayer& l : e->layers) {
        l.a = cur;
        cur += round_up(size_t(max_batch) * l.ld, kAlignFloats);
        l.dz = cur;
        cur += round_up(size_t(max_batch) * l.ld, kAlignFloats);
        if (l.d.type == B200CNN_LAYER_POOL) {
            l.argmax = reinterpret_cast<int32_t*>(cur);
            cur += round_up(size_t(max_batch) * l.nnodes, kAlignFloats);
        }
    }
    if (cudaStreamCreateWithFlags(&e->side, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ev_dz, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ev_data, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ev_side, cudaEventDisableTiming) != cudaSuccess) {
        set_error(int(cudaGetLastError()), "engine_create: cannot create the side stream");
        cudaFree(e->arena);
        delete e;
        return nullptr;
    }
    e->scratch = cur;
    cur += e->scratch_bytes / sizeof(float);
    e->scratch1 = cur;
    cur += e->scratch_bytes / sizeof(float);
    e->metrics = cur;
    e->lr_dev = cur + 8;
    return e;
}

extern "C" void b200cnn_engine_destroy(b200cnn_engine* e) {
    if (!e) return;
    for (auto& ge : e->graphs)
        for (cudaGraphExec_t g : ge.exec)
            if (g) cudaGraphExecDestroy(g);
    if (e->own_buffers) {
        cudaFree(e->params);
        cudaFree(e->grads);
    }
    for (int i = 0; i < 2; i++) {
        if (e->stage_x[i]) cudaFree(e->stage_x[i]);
        if (e->stage_y[i]) cudaFree(e->stage_y[i]);
        if (e->ev_copied[i]) cudaEventDestroy(e->ev_copied[i]);
        if (e->ev_consumed[i]) cudaEventDestroy(e->ev_consumed[i]);
    }
    if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    if (e->ev_dz) cudaEventDestroy(e->ev_dz);
    if (e->ev_data) cudaEventDestroy(e->ev_data);
    if (e->ev_side) cudaEventDestroy(e->ev_side);
    if (e->side) cudaStreamDestroy(e->side);
    cudaFree(e->arena);
    delete e;
}

extern "C" size_t b200cnn_engine_num_params(const b200cnn_engine* e) { return e->dense_params; }
extern "C" size_t b200cnn_engine_flat_floats(const b200cnn_engine* e) { return e->flat_floats; }

extern "C" void b200cnn_engine_param_layout(const b200cnn_engine* e, long* offsets, long* sizes) {
    for (size_t i = 0; i < e->layers.size(); i++) {
        const EngineLayer& l = e->layers[i];
        offsets[2 * i] = l.w_off;
        offsets[2 * i + 1] = l.b_off;
        sizes[2 * i] = l.nw;
        sizes[2 * i + 1] = l.nb;
    }
}

static int ensure_buffers(b200cnn_engine* e) {
    if (e->params) return 0;
    const size_t bytes = e->flat_floats * sizeof(float);
    B200CNN_CUDA(cudaMalloc(&e->params, bytes));
    B200CNN_CUDA(cudaMalloc(&e->grads, bytes));
    B200CNN_CUDA(cudaMemset(e->params, 0, bytes));
    B200CNN_CUDA(cudaMemset(e->grads, 0, bytes));
    e->own_buffers = true;
    return 0;
}

extern "C" int b200cnn_engine_bind_buffers(b200cnn_engine* e, float* params, float* grads) {
    if (e->params) return set_error(kInvalidValue, "engine_bind_buffers: buffers already in use");
    if (!params || !grads || ((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads)) & 127))
        return set_error(kInvalidValue, "engine_bind_buffers: need two 128-byte aligned device buffers");
    e->params = params;
    e->grads = grads;
    e->own_buffers = false;
    return 0;
}

extern "C" float* b200cnn_engine_params(b200cnn_engine* e) { return ensure_buffers(e) ? nullptr : e->params; }
extern "C" float* b200cnn_engine_grads(b200cnn_engine* e) { return ensure_buffers(e) ? nullptr : e->grads; }

static int dense_transfer(b200cnn_engine* e, float* dev_flat, float* dense, bool to_device) {
    if (int rc = ensure_buffers(e)) return rc;
    std::vector<float> flat(e->flat_floats, 0.0f);
    if (!to_device) B200CNN_CUDA(cudaMemcpy(flat.data(), dev_flat, flat.size() * sizeof(float), cudaMemcpyDeviceToHost));
    size_t d = 0;
    for (const EngineLayer& l : e->layers) {
        if (!l.nw) continue;
        if (to_device) {
            std::memcpy(&flat[l.w_off], dense + d, l.nw * sizeof(float));
            std::memcpy(&flat[l.b_off], dense + d + l.nw, l.nb * sizeof(float));
        } else {
            std::memcpy(dense + d, &flat[l.w_off], l.nw * sizeof(float));
            std::memcpy(dense + d + l.nw, &flat[l.b_off], l.nb * sizeof(float));
        }
        d += size_t(l.nw + l.nb);
    }
    if (to_device) B200CNN_CUDA(cudaMemcpy(dev_flat, flat.data(), flat.size() * sizeof(float), cudaMemcpyHostToDevice));
    return 0;
}

extern "C" int b200cnn_engine_set_params(b200cnn_engine* e, const float* dense_host) {
    return dense_transfer(e, e->params ? e->params : b200cnn_engine_params(e), const_cast<float*>(dense_host), true);
}
extern "C" int b200cnn_engine_get_params(b200cnn_engine* e, float* dense_host) {
    if (e->comm) {  // data parallel: parts of the last update are written by the peers; wait until they are here
        if (int rc = b200cnn_comm_wait_weights(e->comm, nullptr)) return rc;
        B200CNN_CUDA(cudaDeviceSynchronize());
    }
    return dense_transfer(e, b200cnn_engine_params(e), dense_host, false);
}
extern "C" int b200cnn_engine_get_grads(b200cnn_engine* e, float* dense_host) {
    if (int rc = ensure_buffers(e)) return rc;
    return dense_transfer(e, cur_grads(e), dense_host, false);
}

extern "C" int b200cnn_engine_forward(b200cnn_engine* e, const void* images, int x_is_u8, const uint8_t* labels,
                                      int batch, b200cnn_stream stream) {
    if (batch < 1 || batch > e->max_batch) return set_error(kInvalidValue, "engine_forward: batch %d outside [1, %d]", batch, e->max_batch);
    if (int rc = ensure_buffers(e)) return rc;
    e->last_images = images;
    e->last_is_u8 = x_is_u8;
    e->last_batch = batch;
    if (e->comm)  // the peers' shards of the last update must have landed before anything reads the parameters
        if (int rc = b200cnn_comm_wait_weights(e->comm, stream)) return rc;
    e->keep_full = true;  // a caller of forward() may look at every layer
    return enqueue_forward(e, images, x_is_u8, labels, batch, as_stream(stream), nullptr);
}

extern "C" int b200cnn_engine_backward(b200cnn_engine* e, b200cnn_stream stream) {
    if (!e->last_batch) return set_error(kInvalidValue, "engine_backward: no forward pass to differentiate");
    b200cnn_comm* saved = e->comm;
    e->comm = nullptr;  // plain backward never communicates
    const int rc = enqueue_backward(e, e->last_images, e->last_is_u8, e->last_batch, as_stream(stream), nullptr);
    e->comm = saved;
    return rc;
}

extern "C" int b200cnn_engine_sgd(b200cnn_engine* e, float lr, b200cnn_stream stream) {
    if (int rc = ensure_buffers(e)) return rc;
    if (int rc = set_lr(e, lr, as_stream(stream))) return rc;
    b200cnn_comm* saved = e->comm;
    e->comm = nullptr;
    const int rc = enqueue_update(e, as_stream(stream), nullptr);
    e->comm = saved;
    return rc;
}

static int enqueue_step(b200cnn_engine* e, const void* images, int is_u8, const uint8_t* labels, int batch,
                        cudaStream_t st, int* launches, bool update = true) {
    const bool chain = use_chain(e, batch);
    const bool mid = !chain && use_mid(e, batch);
    const bool head = !chain && (mid || use_head(e, batch));
    const CriticalStreamScope critical(st);  // kernels on `st` outrank the weight-gradient kernels of the side stream
    e->keep_full = false;  // a training step needs only the pooled activations of a fused conv + pool layer
    if (graph_root_requested()) {
        B200CNN_CUDA(launch_kernel(graph_root_kernel, dim3(1), dim3(32), 0, st));
        if (launches) (*launches)++;
    }
    if (e->comm)
        if (int rc = comm_begin_step(e->comm, st)) return rc;
    if (chain) {
        // layers below the fully connected part, then ONE cluster kernel for fc -> fc -> softmax/loss and all three
        // data-gradient GEMMs (tc_chain.cu)
        auto& L = e->layers;
        const size_t n = L.size();
        if (int rc = enqueue_forward(e, images, is_u8, labels, batch, st, launches, n - 3)) return rc;
        if (e->comm)
            if (int rc = comm_weights_ready(e->comm, st)) return rc;
        EngineLayer &l0 = L[n - 4], &l1 = L[n - 3], &l2 = L[n - 2], &l3 = L[n - 1];
        if (int rc = tc_fc_chain(l0.a, l0.act, e->params + l1.w_off, e->params + l1.b_off, e->params + l2.w_off, e->params + l2.b_off,
                                 e->params + l3.w_off, e->params + l3.b_off, labels, l1.a, l2.a, l3.a, l3.dz, l2.dz, l1.dz, l0.dz,
                                 e->metrics, batch, l0.nnodes, l1.nnodes, l2.nnodes, l3.nnodes, l3.ld, st))
            return rc;
        if (launches) (*launches)++;
    } else if (!head) {
        if (int rc = enqueue_forward(e, images, is_u8, labels, batch, st, launches)) return rc;
    } else {
        // layers below the head, then ONE kernel for fc -> softmax/loss -> both backward-data GEMMs
        auto& L = e->layers;
        const size_t n = L.size();
        if (int rc = enqueue_forward(e, images, is_u8, labels, batch, st, launches, n - 2)) return rc;
        EngineLayer &l1 = L[n - 3], &l2 = L[n - 2], &l3 = L[n - 1];
        if (int rc = (mid ? tc_fc_mid : tc_head)(l1.a, e->params + l2.w_off, e->params + l2.b_off, e->params + l3.w_off, e->params + l3.b_off,
                                                 labels, l2.a, l3.a, l3.dz, l2.dz, l1.dz, e->metrics, batch, l1.nnodes, l2.nnodes, l3.nnodes,
                                                 l3.ld, st))
            return rc;
        if (launches) (*launches)++;
    }
    const bool fuse = update && can_fuse_update(e);
    if (int rc = enqueue_backward(e, images, is_u8, batch, st, launches, chain ? 3 : head ? 2 : 0, fuse)) return rc;
    return (update && !fuse) ? enqueue_update(e, st, launches) : 0;
}

static int train_step_impl(b200cnn_engine* e, const void* images, int x_is_u8, const uint8_t* labels, int batch, float lr,
                           int use_graph, b200cnn_stream stream, bool update);

extern "C" int b200cnn_engine_train_step(b200cnn_engine* e, const void* images, int x_is_u8, const uint8_t* labels,
                                         int batch, float lr, int use_graph, b200cnn_stream stream) {
    return train_step_impl(e, images, x_is_u8, labels, batch, lr, use_graph, stream, true);
}

// forward + backward only (gradients left in the gradient buffer), graph-captured like a full step: the
// building block of update paths that live outside the engine (the NCCL baseline).
extern "C" int b200cnn_engine_forward_backward(b200cnn_engine* e, const void* images, int x_is_u8, const uint8_t* labels,
                                               int batch, int use_graph, b200cnn_stream stream) {
    if (e->comm) return set_error(kInvalidValue, "engine_forward_backward: detach the fused communicator first");
    return train_step_impl(e, images, x_is_u8, labels, batch, 0.0f, use_graph, stream, false);
}

static int train_step_impl(b200cnn_engine* e, const void* images, int x_is_u8, const uint8_t* labels, int batch, float lr,
                           int use_graph, b200cnn_stream stream, bool update) {
    if (batch < 1 || batch > e->max_batch) return set_error(kInvalidValue, "engine_train_step: batch %d outside [1, %d]", batch, e->max_batch);
    if (!labels) return set_error(kInvalidValue, "engine_train_step: labels are required");
    if (int rc = ensure_buffers(e)) return rc;
    cudaStream_t st = as_stream(stream);
    if (update)
        if (int rc = set_lr(e, lr, st)) return rc;
    e->last_images = images;
    e->last_is_u8 = x_is_u8;
    e->last_batch = batch;
    // Data-parallel: this step's gradients go to the other half of the double buffer, so peers that
    // are still reading the previous step's gradients are never overwritten.
    if (e->comm) comm_flip_parity(e->comm);
    const int par = e->comm ? comm_parity(e->comm) : 0;
    if (!use_graph) {
        int launches = 0;
        const int rc = enqueue_step(e, images, x_is_u8, labels, batch, st, &launches, update);
        e->launches_per_step = launches;
        return rc;
    }
    // find (or make room for) the cache entry of this key
    b200cnn_engine::GraphEntry* ge = nullptr;
    for (auto& c : e->graphs)
        if (c.images == images && c.labels == labels && c.batch == batch && c.is_u8 == x_is_u8 && c.stream == st &&
            c.backend == current_backend() && c.update == int(update) && (c.exec[0] || c.exec[1]))
            ge = &c;
    if (!ge) {
        ge = &e->graphs[0];
        for (auto& c : e->graphs)
            if (c.last_use < ge->last_use) ge = &c;  // least recently used
        for (cudaGraphExec_t& g : ge->exec)
            if (g) {
                cudaGraphExecDestroy(g);
                g = nullptr;
            }
        ge->images = images;
        ge->labels = labels;
        ge->batch = batch;
        ge->is_u8 = x_is_u8;
        ge->stream = st;
        ge->backend = current_backend();
        ge->update = int(update);
    }
    ge->last_use = ++e->graph_clock;
    if (!ge->exec[par]) {
        if (st == nullptr) return set_error(kInvalidValue, "engine_train_step: graph capture needs a non-default stream");
        cudaGraph_t graph = nullptr;
        B200CNN_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        int launches = 0;
        const int rc = enqueue_step(e, images, x_is_u8, labels, batch, st, &launches, update);
        const cudaError_t ce = cudaStreamEndCapture(st, &graph);
        if (rc) {
            if (graph) cudaGraphDestroy(graph);
            return rc;
        }
        if (ce != cudaSuccess) return set_error(int(ce), "engine_train_step: capture failed: %s", cudaGetErrorString(ce));
        if (const char* dot = getenv("B200CNN_GRAPH_DOT")) cudaGraphDebugDotPrint(graph, dot, cudaGraphDebugDotFlagsVerbose);  // diagnosis
        const cudaError_t ie = cudaGraphInstantiate(&ge->exec[par], graph, 0);
        cudaGraphDestroy(graph);
        if (ie != cudaSuccess) return set_error(int(ie), "engine_train_step: instantiate failed: %s", cudaGetErrorString(ie));
        e->launches_per_step = launches;
    }
    B200CNN_CUDA(cudaGraphLaunch(ge->exec[par], st));
    return 0;
}

// {sum CE, sum MSE, #correct, #samples} -> mapped pinned host memory (visible to the host once the kernel has completed).
__global__ void store_metrics_to_host_kernel(const float* __restrict__ metrics, float* __restrict__ host_view) {
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x < 4) {
        host_view[threadIdx.x] = __ldcg(metrics + threadIdx.x);
        __threadfence_system();
    }
}

// One training step fed from (pinned) host memory: the step's inputs are copied to a device staging slot on a
// copy stream, the captured step runs on `stream`, and the metrics are copied back to `metrics_host`.  Nothing here
// blocks the host; with two staging slots the copy of batch i+1 overlaps the training of batch i when the caller
// issues steps back to back.  The caller synchronises (stream or event) before reading `metrics_host`.
extern "C" int b200cnn_engine_train_step_host(b200cnn_engine* e, const void* images_host, int x_is_u8, const uint8_t* labels_host,
                                              int batch, float lr, float* metrics_host, b200cnn_stream stream) {
    if (batch < 1 || batch > e->max_batch) return set_error(kInvalidValue, "engine_train_step_host: batch %d outside [1, %d]", batch, e->max_batch);
    cudaStream_t st = as_stream(stream);
    if (st == nullptr) return set_error(kInvalidValue, "engine_train_step_host: needs a non-default stream");
    const EngineLayer& in = e->layers[0];
    const size_t item = size_t(in.nnodes) * (x_is_u8 ? 1 : sizeof(float));
    if (!e->copy_stream) {
        B200CNN_CUDA(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 2; i++) {
            B200CNN_CUDA(cudaEventCreateWithFlags(&e->ev_copied[i], cudaEventDisableTiming));
            B200CNN_CUDA(cudaEventCreateWithFlags(&e->ev_consumed[i], cudaEventDisableTiming));
        }
    }
    if (e->stage_item < item) {  // first call (or a wider input type): (re)allocate the staging slots
        B200CNN_CUDA(cudaDeviceSynchronize());
        for (int i = 0; i < 2; i++) {
            if (e->stage_x[i]) cudaFree(e->stage_x[i]);
            if (e->stage_y[i]) cudaFree(e->stage_y[i]);
            B200CNN_CUDA(cudaMalloc(&e->stage_x[i], size_t(e->max_batch) * item));
            B200CNN_CUDA(cudaMalloc(&e->stage_y[i], size_t(e->max_batch)));
            B200CNN_CUDA(cudaEventRecord(e->ev_consumed[i], st));
        }
        e->stage_item = item;
    }
    const int slot = e->host_slot;
    e->host_slot ^= 1;
    B200CNN_CUDA(cudaStreamWaitEvent(e->copy_stream, e->ev_consumed[slot], 0));  // the step that last read this slot is done
    B200CNN_CUDA(cudaMemcpyAsync(e->stage_x[slot], images_host, size_t(batch) * item, cudaMemcpyHostToDevice, e->copy_stream));
    B200CNN_CUDA(cudaMemcpyAsync(e->stage_y[slot], labels_host, size_t(batch), cudaMemcpyHostToDevice, e->copy_stream));
    B200CNN_CUDA(cudaEventRecord(e->ev_copied[slot], e->copy_stream));
    B200CNN_CUDA(cudaStreamWaitEvent(st, e->ev_copied[slot], 0));
    if (int rc = b200cnn_engine_train_step(e, e->stage_x[slot], x_is_u8, e->stage_y[slot], batch, lr, 1, stream)) return rc;
    B200CNN_CUDA(cudaEventRecord(e->ev_consumed[slot], st));
    if (metrics_host) {
        // The step's result goes back to the host.  A 16-byte DMA between two graph launches costs 3-5 us of stream time;
        // when the destination is pinned memory the device can address (UVA), a four-thread kernel storing straight into
        // it costs less.  B200CNN_E2E_MEMCPY=1 keeps the copy engine.
        static const bool force_memcpy = [] {
            const char* v = getenv("B200CNN_E2E_MEMCPY");
            return v && v[0] == '1';
        }();
        cudaPointerAttributes attr{};
        float* dev_view = nullptr;
        if (!force_memcpy && cudaPointerGetAttributes(&attr, metrics_host) == cudaSuccess && attr.type == cudaMemoryTypeHost && attr.devicePointer)
            dev_view = static_cast<float*>(attr.devicePointer);
        cudaGetLastError();  // (pageable memory: the query may report an error that must not stick)
        if (dev_view)
            B200CNN_CUDA(launch_kernel(store_metrics_to_host_kernel, dim3(1), dim3(32), 0, st, static_cast<const float*>(e->metrics), dev_view));
        else
            B200CNN_CUDA(cudaMemcpyAsync(metrics_host, e->metrics, 4 * sizeof(float), cudaMemcpyDeviceToHost, st));
    }
    return 0;
}

extern "C" float* b200cnn_engine_metrics(b200cnn_engine* e) { return e->metrics; }

extern "C" int b200cnn_engine_reset_metrics(b200cnn_engine* e, b200cnn_stream stream) {
    B200CNN_CUDA(cudaMemsetAsync(e->metrics, 0, 4 * sizeof(float), as_stream(stream)));
    return 0;
}

extern "C" int b200cnn_engine_layer_output(b200cnn_engine* e, int layer, int batch, float* host) {
    if (layer < 1 || layer >= int(e->layers.size()) || batch < 1 || batch > e->max_batch)
        return set_error(kInvalidValue, "engine_layer_output: bad layer / batch");
    const EngineLayer& l = e->layers[size_t(layer)];
    if (layer == 1 && e->l1_stale)
        return set_error(kInvalidValue, "engine_layer_output: a training step does not store the un-pooled activations of the fused conv + pool layer; call forward()");
    B200CNN_CUDA(cudaMemcpy2D(host, l.nnodes * sizeof(float), l.a, l.ld * sizeof(float), l.nnodes * sizeof(float),
                              size_t(batch), cudaMemcpyDeviceToHost));
    if ((e->nhwc_mask >> layer) & 1u) {
        // The tensor-core chain keeps this layer NHWC on the device; callers always get CHW.
        const int c = l.d.depth, hw = l.d.width * l.d.height;
        std::vector<float> tmp(size_t(l.nnodes));
        for (int b = 0; b < batch; b++) {
            float* row = host + size_t(b) * l.nnodes;
            std::memcpy(tmp.data(), row, tmp.size() * sizeof(float));
            for (int p = 0; p < hw; p++)
                for (int ch = 0; ch < c; ch++) row[size_t(ch) * hw + p] = tmp[size_t(p) * c + ch];
        }
    }
    return 0;
}

extern "C" int b200cnn_engine_launches_per_step(const b200cnn_engine* e) { return e->launches_per_step; }
extern "C" int b200cnn_engine_conv_path(const b200cnn_engine* e) {
    return use_fast_chain(e) ? B200CNN_CONV_PATH_TC : use_pool_chain(e) ? B200CNN_CONV_PATH_TC_POOL : B200CNN_CONV_PATH_GENERIC;
}

// Data-parallel hook (declared in comm.cuh): the communicator must have been created over this
// engine's bound parameter / gradient buffers.
extern "C" int b200cnn_engine_attach_comm(b200cnn_engine* e, b200cnn_comm* comm) {
    for (auto& ge : e->graphs)
        for (cudaGraphExec_t& g : ge.exec)
            if (g) {
                cudaGraphExecDestroy(g);
                g = nullptr;
            }
    if (!comm) {
        e->comm = nullptr;
        return 0;
    }
    if (e->params && e->params != b200cnn_comm_params(comm))
        return set_error(kInvalidValue, "engine_attach_comm: attach before the engine allocates its own parameter buffers");
    e->params = b200cnn_comm_params(comm);
    e->grads = b200cnn_comm_grads(comm, 0);
    e->own_buffers = false;
    e->comm = comm;
    // Which layers can a deferred update touch?  Replay the bucketing of layer_grads_ready(): every bucket launched
    // before the last one may be a two-shot push call (peers write this rank's parameters after the call returns);
    // the last bucket is one only if it is too big for the one-shot form.
    e->wait_layer = int(e->layers.size());
    long lo = -1, hi = -1;
    int first_param_layer = 1;
    for (size_t li = 1; li < e->layers.size(); li++)
        if (e->layers[li].nw > 0) {
            first_param_layer = int(li);
            break;
        }
    for (size_t li = e->layers.size() - 1; li >= 1; li--) {
        const EngineLayer& l = e->layers[li];
        if (l.nw == 0) continue;
        const long begin = l.w_off, end = l.b_off + long(round_up(size_t(l.nb), kAlignFloats));
        lo = lo < 0 ? begin : std::min(lo, begin);
        hi = std::max(hi, end);
        const bool last = int(li) == first_param_layer;
        if (!last && hi - lo < kBucketFloats) continue;
        if (comm_update_is_deferred(comm, size_t(hi - lo), last ? 1 : 0)) e->wait_layer = std::min(e->wait_layer, int(li));
        lo = hi = -1;
    }
    return 0;
}
