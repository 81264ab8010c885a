// Native single-GPU trainer (configuration 3; replaces /root/reference/CUDAcnn.cu:416-546, whose only GPU
// code was a never-compiling conv forward).  Pure C++ over the library's C ABI: the data set lives on the
// device as uint8, every step gathers a random mini-batch on the device (cnn.c:455 samples with
// replacement) and replays the engine's CUDA graph.
//
//   cudacnn [--rate 0.1 --nepoch 10 --batch-size 256 --seed 0 --log-every 1000 --synthetic N --synthetic-test N
//            --pool --backend auto|simt|tcgen05 --save F --load F --timing]
//           train-images train-labels test-images test-labels
#include <cuda_runtime.h>

#include <chrono>

#include "gpu_app_common.hpp"

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
    app::Options o;
    if (int rc = app::parse(argc, argv, /*dp=*/false, o)) return rc;
    IdxFile tx, ty, ex, ey;
    if (int rc = app::load_data(o, tx, ty, ex, ey)) return rc;

    CU(cudaSetDevice(0));
    CK(b200cnn_set_backend(o.backend));
    const std::vector<b200cnn_layer_desc> layers = app::topology(o.pool);
    b200cnn_engine* eng = b200cnn_engine_create(layers.data(), int(layers.size()), o.batch);
    if (!eng) {
        fprintf(stderr, "engine_create: %s\n", b200cnn_last_error());
        return 1;
    }
    std::vector<float> p32;
    if (int rc = app::initial_params(o, b200cnn_engine_num_params(eng), p32)) return rc;
    CK(b200cnn_engine_set_params(eng, p32.data()));

    const uint32_t train_size = tx.count(), ntests = ex.count();
    const int item = int(tx.item_size()), batch = o.batch;
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
    const long steps = long(o.nepoch) * std::max<long>(train_size / batch, 1);
    long seen = 0, next_log = 0;
    const auto t0 = std::chrono::steady_clock::now();
    for (long step = 0; step < steps; step++) {
        CK(b200cnn_gather_batch(d_tx, d_ty, train_size, item, xs, ys, batch, o.seed + 1, uint64_t(step), 0, train_size, st));
        CK(b200cnn_engine_train_step(eng, xs, 1, ys, batch, float(o.rate / batch), 1, st));
        seen += batch;
        if (seen > next_log) {
            CU(cudaMemcpyAsync(metrics, b200cnn_engine_metrics(eng), sizeof(metrics), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            CK(b200cnn_engine_reset_metrics(eng, st));
            if (metrics[3] > 0) fprintf(stderr, "i=%ld, error=%.4f\n", seen - batch, metrics[1] / metrics[3]);
            next_log += std::max(o.log_every, batch);
        }
    }
    CU(cudaStreamSynchronize(st));
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (o.timing && steps) fprintf(stderr, "train_images_per_sec=%.1f (%ld samples in %.3f s)\n", steps * batch / dt, steps * batch, dt);
    if (!o.save_path.empty()) {
        CK(b200cnn_engine_get_params(eng, p32.data()));
        if (int rc = app::save_params(o.save_path, p32)) return rc;
    }

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
