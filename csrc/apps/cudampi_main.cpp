// Native multi-GPU data-parallel trainer (configuration 4; replaces /root/reference/CUDAMPI.c + CUDAMPI.cu, which
// needed mpirun + MPI_Init and never selected a GPU).  One process per GPU, forked by this launcher:
//   * rank r owns GPU r; its parameter / gradient region is cudaMalloc'd and shared with the peers through
//     64-byte CUDA IPC handles exchanged over an anonymous shared mapping (no MPI, no NCCL, no Python);
//   * the fused all-reduce + 1/world + SGD kernels of libb200cnn_cuda read the peers' gradients over NVLink;
//   * data is sharded contiguously by rank (cnnmpi.c:456-458), rank 0 evaluates (cnnmpi.c:521).
//
//   cudampi -np N [--rate 0.1 --nepoch 10 --batch-size 256 --seed 0 --log-every 1000 --synthetic M --timing
//                  --algo auto|one_shot|two_shot] train-images train-labels test-images test-labels
#include <cuda_runtime.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../cpu/comm.hpp"
#include "../cpu/idx.hpp"
#include "../cpu/net.hpp"
#include "b200cnn.h"
#include "b200cnn_comm.h"

using namespace b200cnn;

namespace {

struct Options {
    double rate = 0.1;
    int nepoch = 10, batch = 256, log_every = 1000, synthetic = 0, np = 0, algo = B200CNN_ALGO_AUTO;
    unsigned seed = 0;
    bool timing = false;
    std::vector<std::string> pos;
};

struct Shared {  // host-side rendezvous area, mapped before fork()
    unsigned char handles[B200CNN_COMM_MAX_RANKS][64];
    int failed;
};

#define CK(call)                                                                                   \
    do {                                                                                           \
        int rc__ = (call);                                                                         \
        if (rc__) {                                                                                \
            fprintf(stderr, "rank %d: %s failed (%d): %s\n", rank, #call, rc__, b200cnn_last_error()); \
            return 1;                                                                              \
        }                                                                                          \
    } while (0)
#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            fprintf(stderr, "rank %d: %s failed: %s\n", rank, #call, cudaGetErrorString(e__));    \
            return 1;                                                                              \
        }                                                                                          \
    } while (0)

int run_rank(int rank, int world, const Options& o, Shared* sh, ShmComm& host) {
    CU(cudaSetDevice(rank));
    IdxFile tx, ty, ex, ey;
    if (o.synthetic > 0 && o.pos.size() < 4) {
        make_synthetic(uint32_t(o.synthetic), 28, 28, 10, 1, tx, ty);
        make_synthetic(uint32_t(std::max(100, o.synthetic / 5)), 28, 28, 10, 2, ex, ey);
    } else {
        std::string err;
        IdxFile* f[4] = {&tx, &ty, &ex, &ey};
        for (int i = 0; i < 4; i++)
            if (!idx_read(o.pos[i], *f[i], &err)) {
                fprintf(stderr, "%s\n", err.c_str());
                return 111;
            }
    }
    const b200cnn_layer_desc layers[6] = {{B200CNN_LAYER_INPUT, 1, 28, 28, 0, 0, 1},   {B200CNN_LAYER_CONV, 16, 14, 14, 3, 1, 2},
                                          {B200CNN_LAYER_CONV, 32, 7, 7, 3, 1, 2},     {B200CNN_LAYER_FULL, 200, 1, 1, 0, 0, 1},
                                          {B200CNN_LAYER_FULL, 200, 1, 1, 0, 0, 1},    {B200CNN_LAYER_FULL, 10, 1, 1, 0, 0, 1}};
    b200cnn_engine* eng = b200cnn_engine_create(layers, 6, o.batch);
    if (!eng) {
        fprintf(stderr, "rank %d: engine_create: %s\n", rank, b200cnn_last_error());
        return 1;
    }
    // Peer-mapped region: allocate, publish the IPC handle, open everybody else's.
    const size_t flat = b200cnn_engine_flat_floats(eng);
    const size_t bytes = b200cnn_comm_region_bytes(flat);
    void* mine = b200cnn_ipc_alloc(bytes);
    if (!mine) {
        fprintf(stderr, "rank %d: ipc_alloc: %s\n", rank, b200cnn_last_error());
        return 1;
    }
    CK(b200cnn_ipc_get_handle(mine, sh->handles[rank]));
    host.barrier();
    void* regions[B200CNN_COMM_MAX_RANKS] = {};
    for (int r = 0; r < world; r++) {
        regions[r] = (r == rank) ? mine : b200cnn_ipc_open(sh->handles[r]);
        if (!regions[r]) {
            fprintf(stderr, "rank %d: ipc_open(%d): %s\n", rank, r, b200cnn_last_error());
            return 1;
        }
    }
    host.barrier();
    b200cnn_comm* comm = b200cnn_comm_create(rank, world, regions, nullptr, flat);
    if (!comm) {
        fprintf(stderr, "rank %d: comm_create: %s\n", rank, b200cnn_last_error());
        return 1;
    }
    CK(b200cnn_comm_set_algo(comm, o.algo));
    CK(b200cnn_engine_attach_comm(eng, comm));

    // Reference initialisation on every rank, then rank 0's parameters over NVLink (the reference seeds per rank
    // and never synchronises its replicas: SURVEY D7).
    GlibcRand rng(o.seed);
    Net cpu = Net::mnist(rng);
    std::vector<double> p64;
    cpu.gather_params(p64);
    std::vector<float> p32(p64.begin(), p64.end());
    CK(b200cnn_engine_set_params(eng, p32.data()));
    cudaStream_t st;
    CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    CU(cudaDeviceSynchronize());
    host.barrier();
    CK(b200cnn_comm_broadcast_params(comm, 0, st));
    CU(cudaStreamSynchronize(st));
    host.barrier();

    const uint32_t train_size = tx.count(), ntests = ex.count();
    const uint32_t shard = train_size / world, start = shard * rank;
    fprintf(stderr, "%d %u %u\n", rank, start, start + shard);  // cnnmpi.c:459
    const int item = int(tx.item_size());
    uint8_t *d_tx, *d_ty, *d_ex, *d_ey, *xs, *ys;
    CU(cudaMalloc(&d_tx, tx.data.size()));
    CU(cudaMalloc(&d_ty, ty.data.size()));
    CU(cudaMalloc(&d_ex, ex.data.size() + size_t(o.batch) * item));
    CU(cudaMalloc(&d_ey, ey.data.size() + o.batch));
    CU(cudaMalloc(&xs, size_t(o.batch) * item));
    CU(cudaMalloc(&ys, o.batch));
    CU(cudaMemcpy(d_tx, tx.data.data(), tx.data.size(), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_ty, ty.data.data(), ty.data.size(), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_ex, ex.data.data(), ex.data.size(), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_ey, ey.data.data(), ey.data.size(), cudaMemcpyHostToDevice));

    if (rank == 0) fprintf(stderr, "training...\n");
    float metrics[4];
    CK(b200cnn_engine_reset_metrics(eng, st));
    const long steps_per_epoch = std::max<long>(shard / o.batch, 1);
    long seen = 0, next_log = 0, step = 0;
    host.barrier();
    const auto t0 = std::chrono::steady_clock::now();
    for (int epoch = 0; epoch < o.nepoch; epoch++) {
        if (rank == 0) fprintf(stderr, "epoch = %d\n", epoch);
        for (long s = 0; s < steps_per_epoch; s++, step++) {
            CK(b200cnn_gather_batch(d_tx, d_ty, train_size, item, xs, ys, o.batch, o.seed + 1, uint64_t(step) * world + rank, start,
                                    shard, st));
            // lr multiplies the batch-SUM gradient; the fused kernel divides by the world size.
            CK(b200cnn_engine_train_step(eng, xs, 1, ys, o.batch, float(o.rate / o.batch), 1, st));
            seen += o.batch;
            if (rank == 0 && seen > next_log) {
                CU(cudaMemcpyAsync(metrics, b200cnn_engine_metrics(eng), sizeof(metrics), cudaMemcpyDeviceToHost, st));
                CU(cudaStreamSynchronize(st));
                CK(b200cnn_engine_reset_metrics(eng, st));
                if (metrics[3] > 0) fprintf(stderr, "    idx = %ld, error = %f\n", seen - o.batch, metrics[1] / metrics[3]);
                next_log += std::max(o.log_every, o.batch);
            }
        }
    }
    CU(cudaStreamSynchronize(st));
    host.barrier();
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (b200cnn_comm_error(comm)) {
        fprintf(stderr, "rank %d: a fused all-reduce timed out waiting for a peer\n", rank);
        return 1;
    }
    if (rank == 0 && o.timing && step)
        fprintf(stderr, "train_images_per_sec=%.1f (%ld samples, %d GPUs, %.3f s)\n", step * o.batch * world / dt, step * o.batch * world,
                world, dt);
    std::vector<float> fin(b200cnn_engine_num_params(eng));
    CK(b200cnn_engine_get_params(eng, fin.data()));
    double sum = 0;
    for (float v : fin) sum += v;
    fprintf(stderr, "rank %d param_checksum=%.17g\n", rank, sum);
    if (rank == 0) {
        fprintf(stderr, "testing...\n");
        long ncorrect = 0;
        for (uint32_t i = 0; i < ntests; i += o.batch) {
            const int b = int(std::min<uint32_t>(o.batch, ntests - i));
            CK(b200cnn_engine_reset_metrics(eng, st));
            CK(b200cnn_engine_forward(eng, d_ex + size_t(i) * item, 1, d_ey + i, b, st));
            CU(cudaMemcpyAsync(metrics, b200cnn_engine_metrics(eng), sizeof(metrics), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            ncorrect += long(metrics[2]);
        }
        fprintf(stderr, "ntests=%u, ncorrect=%ld\n", ntests, ncorrect);
    }
    host.barrier();  // nobody unmaps a region a peer may still be reading
    b200cnn_engine_destroy(eng);
    b200cnn_comm_destroy(comm);
    for (int r = 0; r < world; r++)
        if (r != rank) b200cnn_ipc_close(regions[r]);
    host.barrier();
    b200cnn_ipc_free(mine);
    return 0;
}

}  // namespace

int main(int argc, char** argv) {
    Options o;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto val = [&]() -> const char* { return i + 1 < argc ? argv[++i] : nullptr; };
        const char* v = nullptr;
        if (a == "--rate") { if (!(v = val())) return 100; o.rate = atof(v); }
        else if (a == "--nepoch") { if (!(v = val())) return 100; o.nepoch = atoi(v); }
        else if (a == "--batch-size") { if (!(v = val())) return 100; o.batch = atoi(v); }
        else if (a == "--seed") { if (!(v = val())) return 100; o.seed = unsigned(strtoul(v, nullptr, 10)); }
        else if (a == "--log-every") { if (!(v = val())) return 100; o.log_every = atoi(v); }
        else if (a == "--synthetic") { if (!(v = val())) return 100; o.synthetic = atoi(v); }
        else if (a == "-np" || a == "--np") { if (!(v = val())) return 100; o.np = atoi(v); }
        else if (a == "--algo") {
            if (!(v = val())) return 100;
            const std::string s = v;
            o.algo = s == "one_shot" ? B200CNN_ALGO_ONE_SHOT : s == "two_shot" ? B200CNN_ALGO_TWO_SHOT : B200CNN_ALGO_AUTO;
        }
        else if (a == "--timing") o.timing = true;
        else if (a.rfind("--", 0) == 0) { fprintf(stderr, "unknown option %s\n", a.c_str()); return 100; }
        else o.pos.push_back(a);
    }
    if (o.synthetic <= 0 && o.pos.size() < 4) return 100;  // cnn.c:412
    if (o.np < 1 || o.np > B200CNN_COMM_MAX_RANKS || o.batch < 1 || o.nepoch < 0) {
        fprintf(stderr, "usage: %s -np N (1..%d) [options] train-images train-labels test-images test-labels\n", argv[0], B200CNN_COMM_MAX_RANKS);
        return 100;
    }
    // Everything shared is mapped before fork(); CUDA is first touched inside the children.
    void* host_region = ShmComm::create_anonymous(o.np, 1024);
    Shared* sh = static_cast<Shared*>(mmap(nullptr, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0));
    if (!host_region || sh == MAP_FAILED) {
        perror("mmap");
        return 111;
    }
    memset(sh, 0, sizeof(Shared));
    std::vector<pid_t> kids;
    for (int r = 0; r < o.np; r++) {
        const pid_t pid = fork();
        if (pid < 0) {
            perror("fork");
            return 111;
        }
        if (pid == 0) {
            ShmComm host(host_region, r, o.np);
            _exit(run_rank(r, o.np, o, sh, host));
        }
        kids.push_back(pid);
    }
    int worst = 0;
    for (pid_t pid : kids) {
        int stat = 0;
        waitpid(pid, &stat, 0);
        const int rc = WIFEXITED(stat) ? WEXITSTATUS(stat) : 128 + WTERMSIG(stat);
        if (rc > worst) worst = rc;
    }
    return worst;
}
