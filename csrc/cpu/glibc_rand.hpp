// Re-implementation of glibc's rand()/srand() (TYPE_3 additive-feedback generator).
//
// The reference seeds with srand(0) and draws both its weight init and its sample order from
// the same rand() stream (/root/reference/cnn.c:413, :46-49, :455).  To be able to diff our
// CPU engine against the unmodified reference binary we need exactly that stream, on any libc,
// and as an object (the reference's is hidden process-global state).
#pragma once
#include <cstdint>

namespace b200cnn {

class GlibcRand {
public:
    static constexpr int kRandMax = 2147483647;

    explicit GlibcRand(unsigned seed = 1) { reseed(seed); }

    void reseed(unsigned seed) {
        // glibc maps seed 0 to 1.
        if (seed == 0) seed = 1;
        int32_t r[34];
        r[0] = static_cast<int32_t>(seed);
        for (int i = 1; i < 31; i++) {
            // 16807 * r[i-1] mod (2^31 - 1) via Schrage so it never overflows 32 bits.
            int32_t hi = r[i - 1] / 127773, lo = r[i - 1] % 127773;
            int32_t w = 16807 * lo - 2836 * hi;
            if (w < 0) w += 2147483647;
            r[i] = w;
        }
        for (int i = 0; i < 31; i++) state_[i] = static_cast<uint32_t>(r[i]);
        f_ = 3;
        r_ = 0;
        // glibc discards the first 310 outputs.
        for (int i = 0; i < 310; i++) next_raw();
    }

    // Same value sequence as glibc rand(): 31 bits, [0, RAND_MAX].
    int next() { return static_cast<int>(next_raw() >> 1); }

    // cnn.c:46 — uniform in [0,1].
    double rnd() { return static_cast<double>(next()) / kRandMax; }
    // cnn.c:49 — Irwin-Hall(4) approximation of N(0,1).
    double nrnd() {
        double a = rnd();
        double b = rnd();
        double c = rnd();
        double d = rnd();
        return (a + b + c + d - 2.0) * 1.724;
    }

private:
    uint32_t next_raw() {
        state_[f_] += state_[r_];
        uint32_t out = state_[f_];
        if (++f_ >= 31) f_ = 0;
        if (++r_ >= 31) r_ = 0;
        return out;
    }
    uint32_t state_[31];
    int f_, r_;
};

}  // namespace b200cnn
