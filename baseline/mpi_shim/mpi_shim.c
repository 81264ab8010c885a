/* See mpi.h.  Shared segment: sense-reversing barrier + one slot of doubles per rank. */
#include "mpi.h"

#include <fcntl.h>
#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#define SLOT 65536
typedef struct {
    volatile unsigned arrived, sense, ready;
    char pad[116];
    double slots[];
} seg_t;

static seg_t* seg;
static int g_rank, g_size;
static unsigned local_sense;
static char seg_name[64];

static int env_int(const char* a, const char* b, int dflt) {
    const char* v = getenv(a);
    if (!v) v = getenv(b);
    return v ? atoi(v) : dflt;
}

static void barrier(void) {
    local_sense ^= 1u;
    if (__sync_add_and_fetch(&seg->arrived, 1) == (unsigned)g_size) {
        seg->arrived = 0;
        __sync_synchronize();
        seg->sense = local_sense;
    } else {
        while (seg->sense != local_sense) sched_yield();
    }
    __sync_synchronize();
}

int MPI_Init(int* argc, char*** argv) {
    (void)argc; (void)argv;
    g_rank = env_int("MPI_SHIM_RANK", "RANK", 0);
    g_size = env_int("MPI_SHIM_SIZE", "WORLD_SIZE", 1);
    const char* port = getenv("MASTER_PORT");
    snprintf(seg_name, sizeof seg_name, "/b200cnn_mpishim_%s", port ? port : "0");
    size_t bytes = sizeof(seg_t) + sizeof(double) * SLOT * (size_t)(g_size + 1);
    int fd = -1;
    if (g_rank == 0) {
        shm_unlink(seg_name);
        fd = shm_open(seg_name, O_CREAT | O_EXCL | O_RDWR, 0600);
        if (fd < 0 || ftruncate(fd, (off_t)bytes) != 0) return 1;
    } else {
        for (int t = 0; t < 60000 && fd < 0; t++) {
            fd = shm_open(seg_name, O_RDWR, 0600);
            struct stat st;
            if (fd >= 0 && (fstat(fd, &st) != 0 || (size_t)st.st_size < bytes)) { close(fd); fd = -1; }
            if (fd < 0) usleep(1000);
        }
        if (fd < 0) return 1;
    }
    seg = mmap(NULL, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (seg == MAP_FAILED) return 1;
    if (g_rank == 0) { seg->arrived = 0; seg->sense = 0; __sync_synchronize(); seg->ready = 0x600dc0de; }
    else while (seg->ready != 0x600dc0de) usleep(200);
    barrier();
    return MPI_SUCCESS;
}
int MPI_Comm_rank(MPI_Comm c, int* r) { (void)c; *r = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm c, int* s) { (void)c; *s = g_size; return MPI_SUCCESS; }

int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm) {
    (void)type; (void)op; (void)comm;
    double* buf = (double*)recvbuf;
    const double* src = (sendbuf == MPI_IN_PLACE) ? buf : (const double*)sendbuf;
    for (int off = 0; off < count; off += SLOT) {
        int len = count - off < SLOT ? count - off : SLOT;
        memcpy(seg->slots + (size_t)g_rank * SLOT, src + off, sizeof(double) * len);
        barrier();
        double* res = seg->slots + (size_t)g_size * SLOT;
        int lo = (int)((long)len * g_rank / g_size), hi = (int)((long)len * (g_rank + 1) / g_size);
        for (int i = lo; i < hi; i++) {
            double s = 0;
            for (int r = 0; r < g_size; r++) s += seg->slots[(size_t)r * SLOT + i];
            res[i] = s;
        }
        barrier();
        memcpy(buf + off, res, sizeof(double) * len);
        barrier();
    }
    return MPI_SUCCESS;
}
int MPI_Finalize(void) {
    barrier();
    if (g_rank == 0) shm_unlink(seg_name);
    return MPI_SUCCESS;
}
