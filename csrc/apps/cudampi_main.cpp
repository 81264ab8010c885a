// Native multi-GPU data-parallel trainer (configuration 4; replaces /root/reference/CUDAMPI.c + CUDAMPI.cu, which
// needed mpirun + MPI_Init and never selected a GPU).  One process per GPU, forked by this launcher:
//   * rank r owns GPU r; its parameter / gradient region is cudaMalloc'd and shared with the peers through
//     64-byte CUDA IPC handles exchanged over an anonymous shared mapping (no MPI, no NCCL, no Python);
//   * the fused all-reduce + 1/world + SGD kernels of libb200cnn_cuda move gradients over NVLink (push protocol for
//     small slices, two-shot for the big bucket).  CUDA IPC memory has no multicast mapping, so the in-switch (NVLS)
//     variant is only reachable through the Python launcher, which allocates symmetric memory;
//   * data is sharded contiguously by rank (cnnmpi.c:456-458), rank 0 evaluates (cnnmpi.c:521);
//   * a rank that fails (bad file, CUDA error, too few GPUs) marks the shared rendezvous area, so its peers leave
//     their barriers with exit code 112 instead of spinning, and the launcher reports the first failure.
//
//   cudampi -np N [--rate 0.1 --nepoch 10 --batch-size 256 --seed 0 --log-every 1000 --synthetic M --synthetic-test M
//                  --pool --backend auto|simt|tcgen05 --save F --load F --timing
//                  --algo auto|one_shot|two_shot|ll_one_shot|ll_two_shot] train-images train-labels test-images test-labels
#include <cuda_runtime.h>

#include <algorithm>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include <chrono>

#include "../cpu/comm.hpp"
#include "gpu_app_common.hpp"

using namespace b200cnn;

namespace {

struct Shared {  // host-side rendezvous area, mapped before fork()
    unsigned char handles[B200CNN_COMM_MAX_RANKS][64];
};

// Any early exit tells the peers (ShmComm::abort): they are, or soon will be, waiting for this rank in a barrier.
#define FAIL(code)      \
    do {                \
        host.abort();   \
        return (code);  \
    } while (0)
#define CK(call)                                                                                   \
    do {                                                                                           \
        int rc__ = (call);                                                                         \
        if (rc__) {                                                                                \
            fprintf(stderr, "rank %d: %s failed (%d): %s\n", rank, #call, rc__, b200cnn_last_error()); \
            FAIL(1);                                                                               \
        }                                                                                          \
    } while (0)
#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            fprintf(stderr, "rank %d: %s failed: %s\n", rank, #call, cudaGetErrorString(e__));    \
            FAIL(1);                                                                               \
        }                                                                                          \
    } while (0)

int run_rank(int rank, int world, const app::Options& o, Shared* sh, ShmComm& host) {
    IdxFile tx, ty, ex, ey;
    if (int rc = app::load_data(o, tx, ty, ex, ey)) FAIL(rc);
    int ngpu = 0;
    if (cudaGetDeviceCount(&ngpu) != cudaSuccess || ngpu < world) {
        if (rank == 0) fprintf(stderr, "cudampi: -np %d needs %d visible GPUs, found %d\n", world, world, ngpu);
        FAIL(1);
    }
    CU(cudaSetDevice(rank));
    if (rank == 0 && world > 1) {
        // NVLink topology as the CUDA runtime sees it from GPU 0 (informational, like the Python launcher's line): peers
        // reachable by load/store, peers with native atomics over the link, best / worst P2P performance rank.
        int reach = 0, atomics = 0, best = 1 << 30, worst = -1;
        for (int d = 1; d < world; d++) {
            int ok = 0, at = 0, perf = 0;
            if (cudaDeviceCanAccessPeer(&ok, 0, d) == cudaSuccess && ok) reach++;
            if (cudaDeviceGetP2PAttribute(&at, cudaDevP2PAttrNativeAtomicSupported, 0, d) == cudaSuccess && at) atomics++;
            if (cudaDeviceGetP2PAttribute(&perf, cudaDevP2PAttrPerformanceRank, 0, d) == cudaSuccess) {
                best = std::min(best, perf);
                worst = std::max(worst, perf);
            }
        }
        cudaGetLastError();  // (an unsupported attribute query must not poison later error checks)
        fprintf(stderr, "topology: %d GPUs, peers reachable from GPU 0: %d/%d, native atomics: %d/%d, P2P performance rank %d..%d\n", world, reach,
                world - 1, atomics, world - 1, worst < 0 ? 0 : best, worst < 0 ? 0 : worst);
    }
    CK(b200cnn_set_backend(o.backend));
    const std::vector<b200cnn_layer_desc> layers = app::topology(o.pool);
    b200cnn_engine* eng = b200cnn_engine_create(layers.data(), int(layers.size()), o.batch);
    if (!eng) {
        fprintf(stderr, "rank %d: engine_create: %s\n", rank, b200cnn_last_error());
        FAIL(1);
    }
    // Peer-mapped region: allocate, publish the IPC handle, open everybody else's.
    const size_t flat = b200cnn_engine_flat_floats(eng);
    const size_t bytes = b200cnn_comm_region_bytes(flat);
    void* mine = b200cnn_ipc_alloc(bytes);
    if (!mine) {
        fprintf(stderr, "rank %d: ipc_alloc: %s\n", rank, b200cnn_last_error());
        FAIL(1);
    }
    CK(b200cnn_ipc_get_handle(mine, sh->handles[rank]));
    host.barrier();
    void* regions[B200CNN_COMM_MAX_RANKS] = {};
    for (int r = 0; r < world; r++) {
        regions[r] = (r == rank) ? mine : b200cnn_ipc_open(sh->handles[r]);
        if (!regions[r]) {
            fprintf(stderr, "rank %d: ipc_open(%d): %s\n", rank, r, b200cnn_last_error());
            FAIL(1);
        }
    }
    host.barrier();
    b200cnn_comm* comm = b200cnn_comm_create(rank, world, regions, nullptr, flat);
    if (!comm) {
        fprintf(stderr, "rank %d: comm_create: %s\n", rank, b200cnn_last_error());
        FAIL(1);
    }
    CK(b200cnn_comm_set_algo(comm, o.algo));
    CK(b200cnn_engine_attach_comm(eng, comm));

    // Reference initialisation (or the checkpoint) on every rank, then rank 0's parameters over NVLink (the reference
    // seeds per rank and never synchronises its replicas: SURVEY D7).
    std::vector<float> p32;
    if (int rc = app::initial_params(o, b200cnn_engine_num_params(eng), p32)) FAIL(rc);
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
            if (seen > next_log) {
                if (rank == 0) {
                    CU(cudaMemcpyAsync(metrics, b200cnn_engine_metrics(eng), sizeof(metrics), cudaMemcpyDeviceToHost, st));
                    CU(cudaStreamSynchronize(st));
                    CK(b200cnn_engine_reset_metrics(eng, st));
                    if (metrics[3] > 0) fprintf(stderr, "    idx = %ld, error = %f\n", seen - o.batch, metrics[1] / metrics[3]);
                }
                // a kernel that gave up waiting for a peer skipped its update: stop instead of training on
                if (b200cnn_comm_error(comm)) {
                    fprintf(stderr, "rank %d: a fused all-reduce timed out waiting for a peer\n", rank);
                    FAIL(1);
                }
                next_log += std::max(o.log_every, o.batch);
            }
        }
    }
    CU(cudaStreamSynchronize(st));
    host.barrier();
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (b200cnn_comm_error(comm)) {
        fprintf(stderr, "rank %d: a fused all-reduce timed out waiting for a peer\n", rank);
        FAIL(1);
    }
    if (rank == 0 && o.timing && step)
        fprintf(stderr, "train_images_per_sec=%.1f (%ld samples, %d GPUs, %.3f s)\n", step * o.batch * world / dt, step * o.batch * world,
                world, dt);
    CK(b200cnn_engine_get_params(eng, p32.data()));  // (waits for the peers' shares of the last update)
    double sum = 0;
    for (float v : p32) sum += v;
    fprintf(stderr, "rank %d param_checksum=%.17g\n", rank, sum);
    if (rank == 0) {
        if (!o.save_path.empty())
            if (int rc = app::save_params(o.save_path, p32)) FAIL(rc);
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
    app::Options o;
    if (int rc = app::parse(argc, argv, /*dp=*/true, o)) return rc;
    if (o.np < 1 || o.np > B200CNN_COMM_MAX_RANKS) {
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
            ShmComm::mark_failed(host_region);
            break;
        }
        if (pid == 0) {
            ShmComm host(host_region, r, o.np);
            _exit(run_rank(r, o.np, o, sh, host));
        }
        kids.push_back(pid);
    }
    // The first rank that ends badly releases the others from their barriers (they exit 112); report its code.
    int first_bad = kids.size() == size_t(o.np) ? 0 : 111;
    for (size_t left = kids.size(); left > 0; left--) {
        int stat = 0;
        if (waitpid(-1, &stat, 0) < 0) break;
        const int rc = WIFEXITED(stat) ? WEXITSTATUS(stat) : 128 + WTERMSIG(stat);
        if (rc != 0) {
            ShmComm::mark_failed(host_region);
            if (first_bad == 0 || first_bad == kCommExit) first_bad = rc;
        }
    }
    return first_bad;
}
