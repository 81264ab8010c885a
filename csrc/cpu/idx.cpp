#include "idx.hpp"

#include <cmath>
#include <cstdio>
#include <memory>

namespace b200cnn {

namespace {
struct FileCloser {
    void operator()(FILE* f) const {
        if (f) fclose(f);
    }
};
using FilePtr = std::unique_ptr<FILE, FileCloser>;

bool fail(std::string* err, const std::string& msg) {
    if (err) *err = msg;
    return false;
}

uint32_t load_be32(const uint8_t* p) {
    return (uint32_t(p[0]) << 24) | (uint32_t(p[1]) << 16) | (uint32_t(p[2]) << 8) | uint32_t(p[3]);
}

// splitmix64: small, seedable, identical everywhere (the data set must not depend on libc).
struct SplitMix {
    uint64_t s;
    uint64_t next() {
        uint64_t z = (s += 0x9e3779b97f4a7c15ull);
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
        return z ^ (z >> 31);
    }
    double uniform() { return (next() >> 11) * (1.0 / 9007199254740992.0); }
};
}  // namespace

bool idx_read(const std::string& path, IdxFile& out, std::string* err) {
    out = IdxFile{};
    FilePtr fp(fopen(path.c_str(), "rb"));
    if (!fp) return fail(err, "cannot open " + path);
    uint8_t header[4];
    if (fread(header, 1, 4, fp.get()) != 4) return fail(err, "short header in " + path);
    if (header[0] != 0 || header[1] != 0) return fail(err, "bad magic in " + path);
    if (header[2] != 0x08) return fail(err, "unsupported element type in " + path);
    if (header[3] < 1) return fail(err, "zero dimensions in " + path);
    const int ndims = header[3];
    std::vector<uint8_t> raw(4 * ndims);
    if (fread(raw.data(), 4, ndims, fp.get()) != size_t(ndims))
        return fail(err, "short dimension table in " + path);
    size_t nbytes = 1;
    out.dims.resize(ndims);
    for (int i = 0; i < ndims; i++) {
        out.dims[i] = load_be32(&raw[4 * i]);
        nbytes *= out.dims[i];
    }
    out.data.resize(nbytes);
    if (nbytes && fread(out.data.data(), 1, nbytes, fp.get()) != nbytes) {
        out = IdxFile{};
        return fail(err, "short payload in " + path);
    }
    return true;
}

bool idx_write(const std::string& path, const IdxFile& f, std::string* err) {
    FilePtr fp(fopen(path.c_str(), "wb"));
    if (!fp) return fail(err, "cannot create " + path);
    const uint8_t header[4] = {0, 0, 0x08, uint8_t(f.dims.size())};
    if (fwrite(header, 1, 4, fp.get()) != 4) return fail(err, "write failed: " + path);
    for (uint32_t d : f.dims) {
        const uint8_t be[4] = {uint8_t(d >> 24), uint8_t(d >> 16), uint8_t(d >> 8), uint8_t(d)};
        if (fwrite(be, 1, 4, fp.get()) != 4) return fail(err, "write failed: " + path);
    }
    if (!f.data.empty() && fwrite(f.data.data(), 1, f.data.size(), fp.get()) != f.data.size())
        return fail(err, "write failed: " + path);
    return true;
}

void make_synthetic(uint32_t n, uint32_t h, uint32_t w, uint32_t nclass, uint64_t seed,
                    IdxFile& images, IdxFile& labels) {
    images.dims = {n, h, w};
    images.data.assign(size_t(n) * h * w, 0);
    labels.dims = {n};
    labels.data.assign(n, 0);

    // Class prototypes: two oriented sinusoidal gratings whose frequency / phase depend on the
    // class, windowed so the border is dark like MNIST.  The prototypes depend only on the
    // class id (not on `seed`) so a train set and a test set with different seeds agree.
    std::vector<double> proto(size_t(nclass) * h * w);
    for (uint32_t c = 0; c < nclass; c++) {
        const double ang = 3.14159265358979323846 * c / nclass;
        const double fx = std::cos(ang) * (1.0 + 0.35 * c), fy = std::sin(ang) * (1.0 + 0.35 * c);
        for (uint32_t y = 0; y < h; y++)
            for (uint32_t x = 0; x < w; x++) {
                const double u = (x + 0.5) / w - 0.5, v = (y + 0.5) / h - 0.5;
                const double win = std::exp(-8.0 * (u * u + v * v));
                const double g = 0.5 + 0.5 * std::sin(6.28318530717958647692 * (fx * u + fy * v) + 0.7 * c);
                proto[(size_t(c) * h + y) * w + x] = win * g;
            }
    }
    SplitMix rng{seed * 0x2545f4914f6cdd1dull + 0x1234567ull};
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t c = uint32_t(rng.next() % nclass);
        labels.data[i] = uint8_t(c);
        const double gain = 0.75 + 0.25 * rng.uniform();
        uint8_t* img = &images.data[size_t(i) * h * w];
        const double* p = &proto[size_t(c) * h * w];
        for (uint32_t j = 0; j < h * w; j++) {
            double v = gain * p[j] + 0.15 * (rng.uniform() - 0.5);
            v = v < 0 ? 0 : (v > 1 ? 1 : v);
            img[j] = uint8_t(v * 255.0 + 0.5);
        }
    }
}

}  // namespace b200cnn
