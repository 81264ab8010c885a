// Data-parallel CPU trainer entrypoint (configuration 2; replaces /root/reference/cnnmpi.c:412-560
// and the `mpirun -np 8 ./cnnmpi ...` line of /root/reference/Makefile:43-44).
//
// Two ways to start it:
//   cnnmpi -np 4 [options] paths...      this process forks 4 ranks over an anonymous shared mapping
//   RANK=r WORLD_SIZE=w cnnmpi paths...  an external launcher (torchrun, a job script) starts the ranks;
//                                        they meet in the POSIX segment /b200cnn_<MASTER_PORT>
//   mpirun -np 4 cnnmpi_mpi paths...     the same program built with -DB200CNN_HAVE_MPI (`make cnnmpi_mpi`): a real
//                                        MPI library carries the all-reduce (csrc/cpu/mpi_comm.hpp)
#include <sys/wait.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "mpi_comm.hpp"
#include "trainer.hpp"

using namespace b200cnn;

namespace {
constexpr size_t kSlotDoubles = 1u << 19;  // 4 MiB staging slot per rank; larger messages are chunked

// Wait for every forked rank.  The first one that ends badly (non-zero exit, signal) releases its peers from whatever
// barrier they are in (they exit with kCommExit); the launcher reports the code of the rank that failed FIRST.
int reap_ranks(const std::vector<pid_t>& kids, void* region) {
    int first_bad = 0;
    for (size_t left = kids.size(); left > 0; left--) {
        int st = 0;
        if (waitpid(-1, &st, 0) < 0) break;
        const int rc = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
        if (rc != 0) {
            ShmComm::mark_failed(region);
            if (first_bad == 0 || first_bad == kCommExit) first_bad = rc;
        }
    }
    return first_bad;
}
}

int main(int argc, char** argv) {
    TrainOptions opt;
    if (int rc = parse_options(argc, argv, /*dp=*/true, opt)) return rc;
#ifdef B200CNN_HAVE_MPI
    // Started by mpirun (or any launcher the MPI library understands), like the reference: `mpirun -np 8 ./cnnmpi ...`
    // (/root/reference/Makefile:44).  Rank and size come from the library; -np is ignored.
    {
        MpiComm comm(&argc, &argv);
        return run_data_parallel(opt, comm);
    }
#endif

    const char* env_rank = getenv("RANK");
    const char* env_world = getenv("WORLD_SIZE");
    if (opt.np <= 0 && env_rank && env_world) {
        const int rank = atoi(env_rank), world = atoi(env_world);
        const char* port = getenv("MASTER_PORT");
        const std::string name = std::string("/b200cnn_") + (port ? port : "0");
        void* region = ShmComm::open_named(name, rank, world, kSlotDoubles);
        if (!region) {
            fprintf(stderr, "rank %d: cannot attach shared segment %s\n", rank, name.c_str());
            return kFileError;
        }
        ShmComm comm(region, rank, world);
        const int rc = run_data_parallel(opt, comm);
        comm.barrier();
        if (rank == 0) ShmComm::unlink_named(name);
        return rc;
    }

    const int world = opt.np > 0 ? opt.np : 1;
    if (world == 1) {
        SelfComm comm;
        return run_data_parallel(opt, comm);
    }
    void* region = ShmComm::create_anonymous(world, kSlotDoubles);
    if (!region) {
        perror("mmap");
        return kFileError;
    }
    std::vector<pid_t> kids;
    for (int r = 0; r < world; r++) {
        const pid_t pid = fork();
        if (pid < 0) {
            perror("fork");
            return kFileError;
        }
        if (pid == 0) {
            ShmComm comm(region, r, world);
            _exit(run_data_parallel(opt, comm));
        }
        kids.push_back(pid);
    }
    return reap_ranks(kids, region);
}
