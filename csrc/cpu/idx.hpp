// IDX (MNIST) container reader / writer and the synthetic MNIST-shaped data generator.
//
// Format as parsed by the reference (/root/reference/cnn.c:352-383): 2 zero bytes, a type byte
// (0x08 = uint8), a dimension-count byte, big-endian uint32 dims, then the raw uint8 payload.
// Unlike three of the reference's four copies the payload is always read, and a
// short read is an error instead of silently training on uninitialised heap.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace b200cnn {

struct IdxFile {
    std::vector<uint32_t> dims;
    std::vector<uint8_t> data;

    uint32_t count() const { return dims.empty() ? 0 : dims[0]; }
    size_t item_size() const {
        size_t n = 1;
        for (size_t i = 1; i < dims.size(); i++) n *= dims[i];
        return n;
    }
    // cnn.c:395 — label i.
    uint8_t get1(size_t i) const { return data[i]; }
    // cnn.c:399 — pointer to image i (dims[1]*dims[2] bytes).
    const uint8_t* get3(size_t i) const { return data.data() + i * item_size(); }
};

// Returns false (and leaves `out` empty) on open / header / short-read failure.
bool idx_read(const std::string& path, IdxFile& out, std::string* err = nullptr);
bool idx_write(const std::string& path, const IdxFile& f, std::string* err = nullptr);

// Synthetic MNIST-shaped set: n images of h*w uint8 and n labels in [0, nclass).
// Each class has a fixed smooth prototype; an image is its class prototype plus noise, so a
// network can actually learn it (random labels would pin the reference metric at 0.09).
void make_synthetic(uint32_t n, uint32_t h, uint32_t w, uint32_t nclass, uint64_t seed,
                    IdxFile& images, IdxFile& labels);

}  // namespace b200cnn
