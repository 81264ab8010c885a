// Real-MPI transport for the CPU data-parallel trainer, compiled only with -DB200CNN_HAVE_MPI (`make cnnmpi_mpi`).
// It uses exactly the five calls the reference makes (/root/reference/cnnmpi.c:419, :421, :422, :490, :558):
// MPI_Init, MPI_Comm_rank, MPI_Comm_size, MPI_Allreduce(MPI_IN_PLACE, MPI_DOUBLE, MPI_SUM) and MPI_Finalize, so it
// also builds against the five-call stand-in under baseline/mpi_shim (how tests/test_cpu_dp.py exercises it where
// no MPI library is installed).  Broadcast and barrier are expressed through the all-reduce: the root contributes
// its data and everybody else zeros; a barrier is an all-reduce of one double.
#pragma once
#ifdef B200CNN_HAVE_MPI
#include <mpi.h>

#include <cstring>
#include <vector>

#include "comm.hpp"

namespace b200cnn {

class MpiComm : public Comm {
public:
    MpiComm(int* argc, char*** argv) {
        MPI_Init(argc, argv);
        MPI_Comm_rank(MPI_COMM_WORLD, &rank_);
        MPI_Comm_size(MPI_COMM_WORLD, &size_);
    }
    ~MpiComm() override { MPI_Finalize(); }
    int rank() const override { return rank_; }
    int size() const override { return size_; }
    void barrier() override {
        double one = 1.0;
        allreduce_sum(&one, 1);
    }
    void allreduce_sum(double* buf, size_t n) override {
        // MPI counts are ints: chunk very long buffers.
        for (size_t off = 0; off < n; off += kChunk) {
            const size_t len = n - off < kChunk ? n - off : kChunk;
            MPI_Allreduce(MPI_IN_PLACE, buf + off, int(len), MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD);
        }
    }
    void broadcast(double* buf, size_t n, int root) override {
        if (rank_ != root) std::memset(buf, 0, n * sizeof(double));
        allreduce_sum(buf, n);
    }

private:
    static constexpr size_t kChunk = size_t(1) << 28;
    int rank_ = 0, size_ = 1;
};

}  // namespace b200cnn
#endif
