// Native single-GPU trainer (configuration 3; replaces /root/reference/CUDAcnn.cu:416-546, whose only GPU
// code was a never-compiling conv forward).  Pure C++ over the library's C ABI: the data set lives on the
// device as uint8, every step gathers a random mini-batch on the device (cnn.c:455 samples with
// replacement) and replays the engine's CUDA graph.
//
//   cudacnn [--rate 0.1 --nepoch 10 --batch-size 256 --seed 0 --log-every 1000 --synthetic N --timing]
//           train-images train-labels test-images test-labels
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../cpu/idx.hpp"
#include "../cpu/net.hpp"
#include "b200cnn.h"

using namespace b200cnn;

#define CK(call)                                                                       \
    do {                                                                               \
        int rc__ = (call);                                                             \
        if (rc__) {                                                                    \
            fprintf(stderr, "%s failed (%d): %s\n", #call, rc__, b200cnn_last_error()); \
            return 1;                                                                  \
        }                                                                              \
    } while (0)
#define CU(call)                                                                   \
    do {                                                                           \
        cudaError_t e__ = (call);                                                  \
        if (e__ != cudaSuccess) {                                                  \
            fprintf(stderr, "%s failed: %s\n", #call, cudaGetErrorString(e__));   \
            return 1;                                                              \
        }                                                                          \
    } while (0)

int main(int argc, char** argv) {
    double rate = 0.1;
    int nepoch = 10, batch = 256, log_every = 1000, synthetic = 0;
    unsigned seed = 0;
    bool timing = false;
    std::vector<std::string> pos;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto val = [&]() -> const char* { return i + 1 < argc ? argv[++i] : nullptr; };
        const char* v = nullptr;
        if (a == "--rate") { if (!(v = val())) return 100; rate = atof(v); }
        else if (a == "--nepoch") { if (!(v = val())) return 100; nepoch = atoi(v); }
        else if (a == "--batch-size") { if (!(v = val())) return 100; batch = atoi(v); }
        else if (a == "--seed") { if (!(v = val())) return 100; seed = unsigned(strtoul(v, nullptr, 10)); }
        else if (a == "--log-every") { if (!(v = val())) return 100; log_every = atoi(v); }
        else if (a == "--synthetic") { if (!(v = val())) return 100; synthetic = atoi(v); }
        else if (a == "--timing") timing = true;
        else if (a.rfind("--", 0) == 0) { fprintf(stderr, "unknown option %s\n", a.c_str()); return 100; }
        else pos.push_back(a);
    }
    if (synthetic <= 0 && pos.size() < 4) return 100;  // cnn.c:412
    if (batch < 1 || nepoch < 0) return 100;

    IdxFile tx, ty, ex, ey;
    if (synthetic > 0 && pos.size() < 4) {
        make_synthetic(uint32_t(synthetic), 28, 28, 10, 1, tx, ty);
        make_synthetic(uint32_t(std::max(100, synthetic / 5)), 28, 28, 10, 2, ex, ey);
    } else {
        std::string err;
        IdxFile* f[4] = {&tx, &ty, &ex, &ey};
        for (int i = 0; i < 4; i++)
            if (!idx_read(pos[i], *f[i], &err)) {
                fprintf(stderr, "%s\n", err.c_str());
                return 111;  // cnn.c:432-442
            }
    }

    // Reference topology and initialisation (cnn.c:413-428) through the CPU engine's generator.
    GlibcRand rng(seed);
    Net cpu = Net::mnist(rng);
    std::vector<double> p64;
    cpu.gather_params(p64);
    std::vector<float> p32(p64.begin(), p64.end());
    const b200cnn_layer_desc layers[6] = {{B200CNN_LAYER_INPUT, 1, 28, 28, 0, 0, 1},   {B200CNN_LAYER_CONV, 16, 14, 14, 3, 1, 2},
                                          {B200CNN_LAYER_CONV, 32, 7, 7, 3, 1, 2},     {B200CNN_LAYER_FULL, 200, 1, 1, 0, 0, 1},
                                          {B200CNN_LAYER_FULL, 200, 1, 1, 0, 0, 1},    {B200CNN_LAYER_FULL, 10, 1, 1, 0, 0, 1}};
    CU(cudaSetDevice(0));
    b200cnn_engine* eng = b200cnn_engine_create(layers, 6, batch);
    if (!eng) {
        fprintf(stderr, "engine_create: %s\n", b200cnn_last_error());
        return 1;
    }
    CK(b200cnn_engine_set_params(eng, p32.data()));

    const uint32_t train_size = tx.count(), ntests = ex.count();
    const int item = int(tx.item_size());
    uint8_t *d_tx, *d_ty, *d_ex, *d_ey, *xs, *ys;
    CU(cudaMalloc(&d_tx, tx.data.size()));
    CU(cudaMalloc(&d_ty, ty.data.size()));
    CU(cudaMalloc(&d_ex, ex.data.size() + size_t(batch) * item));
    CU(cudaMalloc(&d_ey, ey.data.size() + batch));
    CU(cudaMalloc(&xs, size_t(batch) * item));
    CU(cudaMalloc(&ys, batch));
    CU(cudaMemcpy(d_tx, tx.data.data(), tx.data.size(), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_ty, ty.data.data(), ty.data.size(), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_ex, ex.data.data(), ex.data.size(), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_ey, ey.data.data(), ey.data.size(), cudaMemcpyHostToDevice));
    cudaStream_t st;
    CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));

    fprintf(stderr, "training...\n");
    float metrics[4];
    CK(b200cnn_engine_reset_metrics(eng, st));
    const long steps = long(nepoch) * std::max<long>(train_size / batch, 1);
    long seen = 0, next_log = 0;
    const auto t0 = std::chrono::steady_clock::now();
    for (long step = 0; step < steps; step++) {
        CK(b200cnn_gather_batch(d_tx, d_ty, train_size, item, xs, ys, batch, seed + 1, uint64_t(step), 0, train_size, st));
        CK(b200cnn_engine_train_step(eng, xs, 1, ys, batch, float(rate / batch), 1, st));
        seen += batch;
        if (seen > next_log) {
            CU(cudaMemcpyAsync(metrics, b200cnn_engine_metrics(eng), sizeof(metrics), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            CK(b200cnn_engine_reset_metrics(eng, st));
            if (metrics[3] > 0) fprintf(stderr, "i=%ld, error=%.4f\n", seen - batch, metrics[1] / metrics[3]);
            next_log += std::max(log_every, batch);
        }
    }
    CU(cudaStreamSynchronize(st));
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (timing && steps) fprintf(stderr, "train_images_per_sec=%.1f (%ld samples in %.3f s)\n", steps * batch / dt, steps * batch, dt);

    fprintf(stderr, "testing...\n");
    long ncorrect = 0;
    for (uint32_t i = 0; i < ntests; i += batch) {
        const int b = int(std::min<uint32_t>(batch, ntests - i));
        CK(b200cnn_engine_reset_metrics(eng, st));
        CK(b200cnn_engine_forward(eng, d_ex + size_t(i) * item, 1, d_ey + i, b, st));
        CU(cudaMemcpyAsync(metrics, b200cnn_engine_metrics(eng), sizeof(metrics), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        ncorrect += long(metrics[2]);
        if ((i % 1000) < uint32_t(batch)) fprintf(stderr, "i=%u\n", i);
    }
    fprintf(stderr, "ntests=%u, ncorrect=%ld\n", ntests, ncorrect);
    b200cnn_engine_destroy(eng);
    cudaFree(d_tx); cudaFree(d_ty); cudaFree(d_ex); cudaFree(d_ey); cudaFree(xs); cudaFree(ys);
    cudaStreamDestroy(st);
    return 0;
}
