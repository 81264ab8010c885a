// Host-side communicator for the CPU data-parallel trainer.
//
// The reference talks to an external MPI library through five calls: MPI_Init, MPI_Comm_rank,
// MPI_Comm_size, MPI_Allreduce(IN_PLACE, DOUBLE, SUM) and MPI_Finalize
// (/root/reference/cnnmpi.c:419-422, :490, :558).  No MPI exists in the target image, so the
// transport here is a shared-memory segment between processes of one host: a sense-reversing
// barrier, one staging slot per rank and a reduce-scatter + all-gather sum whose per-element
// addition order is fixed by rank index (bitwise identical result on every rank, every run).
// A real MPI backend (MpiComm, the same five calls) is compiled in with -DB200CNN_HAVE_MPI: `make cnnmpi_mpi`.
//
// Failure handling (the reference has none: a rank that returns 111 before MPI_Finalize leaves the others hanging,
// cnnmpi.c:443): every wait inside ShmComm is bounded and watches a shared `failed` word.  A rank that gives up
// calls abort() (or its launcher does, mark_failed()); every other rank then leaves its barrier and exits with
// kCommExit instead of spinning forever.
#pragma once
#include <atomic>
#include <cstddef>
#include <cstdint>
#include <string>

namespace b200cnn {

class Comm {
public:
    virtual ~Comm() {}
    virtual int rank() const = 0;
    virtual int size() const = 0;
    virtual void barrier() = 0;
    virtual void allreduce_sum(double* buf, size_t n) = 0;  // in place, like MPI_IN_PLACE
    virtual void broadcast(double* buf, size_t n, int root) = 0;
    virtual void abort() {}  // tell the peers this rank is leaving early
};

constexpr int kCommExit = 112;  // exit code of a rank whose peer failed or never arrived

// Trivial single-process communicator.
class SelfComm : public Comm {
public:
    int rank() const override { return 0; }
    int size() const override { return 1; }
    void barrier() override {}
    void allreduce_sum(double*, size_t) override {}
    void broadcast(double*, size_t, int) override {}
};

struct ShmHeader {
    std::atomic<uint32_t> arrived;
    std::atomic<uint32_t> sense;
    std::atomic<uint32_t> attached;
    uint32_t world;
    uint64_t slot_doubles;
    std::atomic<uint32_t> failed;  // set by a rank (or the launcher) that gives up: barriers stop waiting
    std::atomic<uint32_t> joined;  // named segments: non-root ranks that have attached to THIS segment
    std::atomic<uint32_t> go;      // named segments: set by rank 0 once everybody has joined
};

class ShmComm : public Comm {
public:
    // Bytes needed for `world` ranks with staging slots of `slot_doubles` doubles each.
    static size_t region_bytes(int world, size_t slot_doubles);
    // Anonymous shared mapping, to be created BEFORE fork() by a launcher process.
    static void* create_anonymous(int world, size_t slot_doubles);
    // Named POSIX segment for externally launched ranks (RANK / WORLD_SIZE in the environment);
    // rank 0 creates and initialises it, the others wait for it to appear.
    static void* open_named(const std::string& name, int rank, int world, size_t slot_doubles);
    static void unlink_named(const std::string& name);

    ShmComm(void* region, int rank, int world);
    int rank() const override { return rank_; }
    int size() const override { return world_; }
    void barrier() override;
    void allreduce_sum(double* buf, size_t n) override;
    void broadcast(double* buf, size_t n, int root) override;
    void abort() override { hdr_->failed.store(1, std::memory_order_release); }
    // For a launcher that holds the mapping but is not a rank: release everybody who is waiting.
    static void mark_failed(void* region) { static_cast<ShmHeader*>(region)->failed.store(1, std::memory_order_release); }

private:
    double* slot(int r) const { return slots_ + size_t(r) * slot_doubles_; }
    ShmHeader* hdr_;
    double* slots_;   // world slots followed by one result slot
    size_t slot_doubles_;
    int rank_, world_;
    uint32_t local_sense_ = 0;
};

}  // namespace b200cnn
