#include "comm.hpp"

#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

namespace b200cnn {

namespace {
constexpr size_t kHeaderBytes = 256;  // keeps the slots cache-line aligned

double now_seconds() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// Longest a rank waits for its peers in one barrier (B200CNN_COMM_TIMEOUT_S, default 300 s).
double wait_limit_seconds() {
    static const double v = [] {
        const char* e = getenv("B200CNN_COMM_TIMEOUT_S");
        const double d = e ? atof(e) : 0.0;
        return d > 0 ? d : 300.0;
    }();
    return v;
}

// Inode of the segment `name` currently refers to (0 if it does not exist).
ino_t named_inode(const std::string& name) {
    const int fd = shm_open(name.c_str(), O_RDWR, 0600);
    if (fd < 0) return 0;
    struct stat st;
    const ino_t ino = fstat(fd, &st) == 0 ? st.st_ino : 0;
    close(fd);
    return ino;
}

void init_header(void* region, int world, size_t slot_doubles) {
    ShmHeader* h = new (region) ShmHeader;
    h->arrived.store(0);
    h->sense.store(0);
    h->world = uint32_t(world);
    h->slot_doubles = slot_doubles;
    h->failed.store(0);
    h->joined.store(0);
    h->go.store(0);
    h->attached.store(0x600dc0deu, std::memory_order_release);  // "initialised" marker
}
}  // namespace

size_t ShmComm::region_bytes(int world, size_t slot_doubles) {
    return kHeaderBytes + sizeof(double) * slot_doubles * size_t(world + 1);
}

void* ShmComm::create_anonymous(int world, size_t slot_doubles) {
    const size_t bytes = region_bytes(world, slot_doubles);
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (p == MAP_FAILED) return nullptr;
    init_header(p, world, slot_doubles);
    return p;
}

void* ShmComm::open_named(const std::string& name, int rank, int world, size_t slot_doubles) {
    const size_t bytes = region_bytes(world, slot_doubles);
    int fd = -1;
    if (rank == 0) {
        shm_unlink(name.c_str());
        fd = shm_open(name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
        if (fd < 0 || ftruncate(fd, off_t(bytes)) != 0) return nullptr;
    } else {
        // Wait (up to ~60 s) for rank 0 to create and size the segment.
        for (int tries = 0; tries < 60000; tries++) {
            fd = shm_open(name.c_str(), O_RDWR, 0600);
            if (fd >= 0) {
                struct stat st;
                if (fstat(fd, &st) == 0 && size_t(st.st_size) >= bytes) break;
                close(fd);
                fd = -1;
            }
            usleep(1000);
        }
        if (fd < 0) return nullptr;
    }
    struct stat mine;
    if (fstat(fd, &mine) != 0) mine.st_ino = 0;
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) return nullptr;
    ShmHeader* h = static_cast<ShmHeader*>(p);
    const double t0 = now_seconds();
    if (rank == 0) {
        init_header(p, world, slot_doubles);
        // Handshake: everybody must have attached to THIS segment (not to one a crashed run left behind under the
        // same name) before anyone enters a barrier.
        while (h->joined.load(std::memory_order_acquire) != uint32_t(world - 1)) {
            if (now_seconds() - t0 > wait_limit_seconds()) {
                munmap(p, bytes);
                return nullptr;
            }
            usleep(200);
        }
        h->go.store(1, std::memory_order_release);
        return p;
    }
    // A segment whose `go` is already set before this rank has joined belongs to an earlier run (rank 0 raises it only
    // after counting every rank of ITS run); so does one that the name no longer points to.  Drop it and look again.
    bool stale = false;
    while (h->attached.load(std::memory_order_acquire) != 0x600dc0deu && !stale) {
        if (now_seconds() - t0 > 2.0 && named_inode(name) != mine.st_ino) stale = true;
        if (now_seconds() - t0 > wait_limit_seconds()) break;
        usleep(200);
    }
    if (!stale && h->attached.load(std::memory_order_acquire) == 0x600dc0deu) {
        if (h->go.load(std::memory_order_acquire) != 0) {
            stale = true;
        } else {
            h->joined.fetch_add(1, std::memory_order_acq_rel);
            double last_check = now_seconds();
            while (h->go.load(std::memory_order_acquire) == 0) {
                if (now_seconds() - last_check > 1.0) {
                    last_check = now_seconds();
                    if (named_inode(name) != mine.st_ino) {
                        stale = true;
                        break;
                    }
                    if (last_check - t0 > wait_limit_seconds()) break;
                }
                usleep(200);
            }
            if (!stale && h->go.load(std::memory_order_acquire) != 0) return p;
        }
    }
    munmap(p, bytes);
    if (!stale || now_seconds() - t0 > wait_limit_seconds()) return nullptr;
    usleep(20000);
    return open_named(name, rank, world, slot_doubles);  // rank 0 of THIS run replaces the stale segment; try again
}

void ShmComm::unlink_named(const std::string& name) { shm_unlink(name.c_str()); }

ShmComm::ShmComm(void* region, int rank, int world)
    : hdr_(static_cast<ShmHeader*>(region)),
      slots_(reinterpret_cast<double*>(static_cast<char*>(region) + kHeaderBytes)),
      slot_doubles_(hdr_->slot_doubles),
      rank_(rank),
      world_(world) {}

void ShmComm::barrier() {
    local_sense_ ^= 1u;
    if (hdr_->arrived.fetch_add(1, std::memory_order_acq_rel) + 1 == uint32_t(world_)) {
        hdr_->arrived.store(0, std::memory_order_relaxed);
        hdr_->sense.store(local_sense_, std::memory_order_release);
    } else {
        int spins = 0;
        double t0 = 0;
        while (hdr_->sense.load(std::memory_order_acquire) != local_sense_) {
            if (++spins <= 64) continue;
            sched_yield();  // ranks may outnumber cores (mpirun -np 8 on a laptop)
            if ((spins & 1023) != 0) continue;
            // A peer that left early (bad file, CUDA error, crash) will never arrive: stop waiting.
            const bool peer_failed = hdr_->failed.load(std::memory_order_acquire) != 0;
            if (t0 == 0) t0 = now_seconds();
            if (peer_failed || now_seconds() - t0 > wait_limit_seconds()) {
                fprintf(stderr, "rank %d: %s; leaving\n", rank_, peer_failed ? "a peer failed" : "timed out waiting for the other ranks in a barrier");
                hdr_->failed.store(1, std::memory_order_release);
                _exit(kCommExit);
            }
        }
    }
}

void ShmComm::allreduce_sum(double* buf, size_t n) {
    double* result = slot(world_);
    for (size_t off = 0; off < n; off += slot_doubles_) {
        const size_t len = std::min(slot_doubles_, n - off);
        std::memcpy(slot(rank_), buf + off, len * sizeof(double));
        barrier();
        // Reduce-scatter: this rank owns one contiguous slice and sums it over ranks 0..W-1.
        const size_t lo = len * size_t(rank_) / world_, hi = len * size_t(rank_ + 1) / world_;
        for (size_t i = lo; i < hi; i++) {
            double s = slot(0)[i];
            for (int r = 1; r < world_; r++) s += slot(r)[i];
            result[i] = s;
        }
        barrier();
        // All-gather: everyone reads the complete result.
        std::memcpy(buf + off, result, len * sizeof(double));
        barrier();
    }
}

void ShmComm::broadcast(double* buf, size_t n, int root) {
    double* result = slot(world_);
    for (size_t off = 0; off < n; off += slot_doubles_) {
        const size_t len = std::min(slot_doubles_, n - off);
        if (rank_ == root) std::memcpy(result, buf + off, len * sizeof(double));
        barrier();
        if (rank_ != root) std::memcpy(buf + off, result, len * sizeof(double));
        barrier();
    }
}

}  // namespace b200cnn
