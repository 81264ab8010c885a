// Data-parallel CPU trainer entrypoint (configuration 2; replaces /root/reference/cnnmpi.c:412-560
// and the `mpirun -np 8 ./cnnmpi ...` line of /root/reference/Makefile:43-44).
//
// Two ways to start it:
//   cnnmpi -np 4 [options] paths...      this process forks 4 ranks over an anonymous shared mapping
//   RANK=r WORLD_SIZE=w cnnmpi paths...  an external launcher (torchrun, a job script) starts the ranks;
//                                        they meet in the POSIX segment /b200cnn_<MASTER_PORT>
#include <sys/wait.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "trainer.hpp"

using namespace b200cnn;

namespace {
constexpr size_t kSlotDoubles = 1u << 19;  // 4 MiB staging slot per rank; larger messages are chunked
}

int main(int argc, char** argv) {
    TrainOptions opt;
    if (int rc = parse_options(argc, argv, /*dp=*/true, opt)) return rc;

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
    int worst = 0;
    for (pid_t pid : kids) {
        int st = 0;
        waitpid(pid, &st, 0);
        const int rc = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
        if (rc > worst) worst = rc;
    }
    return worst;
}
