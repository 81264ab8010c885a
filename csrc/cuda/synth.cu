// Device-side synthetic MNIST-shaped data (K10): class prototype + noise, counter-based RNG, so the
// single- and multi-GPU trainers can draw disjoint shards without touching the host.
// Reference being replaced: the host-side sample fetch + /255 + one-hot of /root/reference/cnn.c:455-464
// (the /255 and one-hot parts are fused into the first conv and the loss kernel instead).
#include "common.cuh"

namespace b200cnn {
namespace {

__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

__global__ void synthetic_kernel(uint8_t* __restrict__ images, uint8_t* __restrict__ labels, int n, int h, int w,
                                 int nclass, uint64_t seed, uint64_t first) {
    const int img = blockIdx.x;
    if (img >= n) return;
    const uint64_t id = first + img;
    const uint64_t base = mix64(seed * 0x2545f4914f6cdd1dull + id);
    const int c = int(base % uint64_t(nclass));
    const float gain = 0.75f + 0.25f * float((base >> 20) & 0xffff) * (1.0f / 65536.0f);
    if (threadIdx.x == 0) labels[img] = uint8_t(c);
    const float ang = 3.14159265f * c / nclass;
    const float fx = cosf(ang) * (1.0f + 0.35f * c), fy = sinf(ang) * (1.0f + 0.35f * c);
    for (int j = threadIdx.x; j < h * w; j += blockDim.x) {
        const int x = j % w, y = j / w;
        const float u = (x + 0.5f) / w - 0.5f, v = (y + 0.5f) / h - 0.5f;
        const float win = __expf(-8.0f * (u * u + v * v));
        const float g = 0.5f + 0.5f * __sinf(6.2831853f * (fx * u + fy * v) + 0.7f * c);
        const float noise = float(mix64(base ^ (uint64_t(j) * 0x9e3779b97f4a7c15ull)) >> 40) * (1.0f / 16777216.0f) - 0.5f;
        float val = gain * win * g + 0.15f * noise;
        val = fminf(fmaxf(val, 0.0f), 1.0f);
        images[size_t(img) * h * w + j] = uint8_t(val * 255.0f + 0.5f);
    }
}

__global__ void gather_batch_kernel(const uint8_t* __restrict__ images, const uint8_t* __restrict__ labels,
                                    int item_bytes, uint8_t* __restrict__ out_images, uint8_t* __restrict__ out_labels,
                                    int batch, uint64_t seed, uint64_t step, uint32_t shard_begin, uint32_t shard_len) {
    const int b = blockIdx.x;
    if (b >= batch) return;
    pdl_wait();
    pdl_launch_dependents();
    const uint32_t idx = shard_begin + uint32_t(mix64(mix64(seed ^ (step * 0x9e3779b97f4a7c15ull)) + b) % shard_len);
    if (threadIdx.x == 0) out_labels[b] = labels[idx];
    const uint8_t* src = images + size_t(idx) * item_bytes;
    uint8_t* dst = out_images + size_t(b) * item_bytes;
    if (item_bytes % 16 == 0) {
        const uint4* s4 = reinterpret_cast<const uint4*>(src);
        uint4* d4 = reinterpret_cast<uint4*>(dst);
        for (int i = threadIdx.x; i < item_bytes / 16; i += blockDim.x) d4[i] = __ldg(s4 + i);
    } else {
        for (int i = threadIdx.x; i < item_bytes; i += blockDim.x) dst[i] = __ldg(src + i);
    }
}

}  // namespace
}  // namespace b200cnn

using namespace b200cnn;

extern "C" int b200cnn_gather_batch(const uint8_t* images, const uint8_t* labels, uint32_t n_items, int item_bytes,
                                    uint8_t* out_images, uint8_t* out_labels, int batch, uint64_t seed, uint64_t step,
                                    uint32_t shard_begin, uint32_t shard_len, b200cnn_stream stream) {
    if (batch <= 0 || item_bytes <= 0 || shard_len == 0 || shard_begin + shard_len > n_items)
        return set_error(kInvalidValue, "gather_batch: bad shard [%u, %u) of %u", shard_begin, shard_begin + shard_len, n_items);
    B200CNN_CUDA(launch_kernel(gather_batch_kernel, dim3(batch), dim3(64), 0, as_stream(stream), images, labels, item_bytes,
                               out_images, out_labels, batch, seed, step, shard_begin, shard_len));
    return 0;
}

extern "C" int b200cnn_synthetic_batch(uint8_t* images, uint8_t* labels, int n, int h, int w, int nclass,
                                       uint64_t seed, uint64_t first_index, b200cnn_stream stream) {
    if (n <= 0 || h <= 0 || w <= 0 || nclass <= 0 || nclass > 255) return set_error(kInvalidValue, "synthetic_batch: bad shape");
    synthetic_kernel<<<n, 128, 0, as_stream(stream)>>>(images, labels, n, h, w, nclass, seed, first_index);
    B200CNN_CHECK_LAUNCH("synthetic_batch");
    return 0;
}
