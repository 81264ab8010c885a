// Shared by the two native GPU trainers (cudacnn, cudampi): command line, data, topology, initialisation,
// checkpoints.  Same flags as the Python entrypoints (mpi_cuda_cnn_b200/cli.py) so either can be used:
//   --rate --nepoch --batch-size --seed --log-every --synthetic N --synthetic-test N --pool --backend auto|simt|tcgen05
//   --save F --load F --timing, and for cudampi: -np N --algo auto|one_shot|two_shot|ll_one_shot|ll_two_shot
// Defaults are the reference's literals (/root/reference/cnn.c:413, :446-449) except the batch (256 per GPU, the
// benchmark configuration).  Checkpoints are the dense float32 parameters in layer order (weights then biases per
// layer), the format `cli.py --save` writes, so files move freely between the native and the Python programs.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../cpu/idx.hpp"
#include "../cpu/net.hpp"
#include "b200cnn.h"
#include "b200cnn_comm.h"

namespace b200cnn {
namespace app {

struct Options {
    double rate = 0.1;
    int nepoch = 10, batch = 256, log_every = 1000, synthetic = 0, synthetic_test = 0, np = 0;
    int algo = B200CNN_ALGO_AUTO, backend = B200CNN_BACKEND_AUTO;
    unsigned seed = 0;
    bool timing = false, pool = false;
    std::string save_path, load_path;
    std::vector<std::string> pos;
};

// 0 = ok, 100 = usage error (the reference's exit code for a bad command line, cnn.c:412).
inline int parse(int argc, char** argv, bool dp, Options& o) {
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto val = [&]() -> const char* { return i + 1 < argc ? argv[++i] : nullptr; };
        const char* v = nullptr;
        if (a == "--rate") { if (!(v = val())) return 100; o.rate = atof(v); }
        else if (a == "--nepoch") { if (!(v = val())) return 100; o.nepoch = atoi(v); }
        else if (a == "--batch-size") { if (!(v = val())) return 100; o.batch = atoi(v); }
        else if (a == "--seed") { if (!(v = val())) return 100; o.seed = unsigned(strtoul(v, nullptr, 10)); }
        else if (a == "--log-every") { if (!(v = val())) return 100; o.log_every = atoi(v); }
        else if (a == "--synthetic") { if (!(v = val())) return 100; o.synthetic = atoi(v); }
        else if (a == "--synthetic-test") { if (!(v = val())) return 100; o.synthetic_test = atoi(v); }
        else if (a == "--save") { if (!(v = val())) return 100; o.save_path = v; }
        else if (a == "--load") { if (!(v = val())) return 100; o.load_path = v; }
        else if (a == "--pool") o.pool = true;
        else if (a == "--timing") o.timing = true;
        else if (a == "--backend") {
            if (!(v = val())) return 100;
            const std::string s = v;
            if (s == "auto") o.backend = B200CNN_BACKEND_AUTO;
            else if (s == "simt") o.backend = B200CNN_BACKEND_SIMT;
            else if (s == "tcgen05") o.backend = B200CNN_BACKEND_TCGEN05;
            else { fprintf(stderr, "unknown backend %s\n", v); return 100; }
        }
        else if (dp && (a == "-np" || a == "--np")) { if (!(v = val())) return 100; o.np = atoi(v); }
        else if (dp && a == "--algo") {
            if (!(v = val())) return 100;
            const std::string s = v;
            if (s == "auto") o.algo = B200CNN_ALGO_AUTO;
            else if (s == "one_shot") o.algo = B200CNN_ALGO_ONE_SHOT;
            else if (s == "two_shot") o.algo = B200CNN_ALGO_TWO_SHOT;
            else if (s == "ll_one_shot") o.algo = B200CNN_ALGO_LL_ONE_SHOT;
            else if (s == "ll_two_shot") o.algo = B200CNN_ALGO_LL_TWO_SHOT;
            else { fprintf(stderr, "unknown algorithm %s (nvls needs the Python launcher: it maps the multicast address)\n", v); return 100; }
        }
        else if (a.rfind("--", 0) == 0) { fprintf(stderr, "unknown option %s\n", a.c_str()); return 100; }
        else o.pos.push_back(a);
    }
    if (o.synthetic <= 0 && o.pos.size() < 4) return 100;  // cnn.c:412
    if (o.batch < 1 || o.nepoch < 0 || o.log_every < 1) return 100;
    return 0;
}

// 0 = ok, 111 = file error (cnn.c:432-442).
inline int load_data(const Options& o, IdxFile& tx, IdxFile& ty, IdxFile& ex, IdxFile& ey) {
    if (o.synthetic > 0 && o.pos.size() < 4) {
        make_synthetic(uint32_t(o.synthetic), 28, 28, 10, 1, tx, ty);
        make_synthetic(uint32_t(o.synthetic_test > 0 ? o.synthetic_test : std::max(100, o.synthetic / 5)), 28, 28, 10, 2, ex, ey);
        return 0;
    }
    std::string err;
    IdxFile* f[4] = {&tx, &ty, &ex, &ey};
    for (int i = 0; i < 4; i++)
        if (!idx_read(o.pos[i], *f[i], &err)) {
            fprintf(stderr, "%s\n", err.c_str());
            return 111;
        }
    if (tx.dims.size() != 3 || ex.dims.size() != 3 || tx.item_size() != 28 * 28 || ex.item_size() != 28 * 28 || tx.count() != ty.count() ||
        ex.count() != ey.count()) {
        fprintf(stderr, "expected 28x28 images (idx3) with matching 1-D label files\n");
        return 111;
    }
    return 0;
}

// The reference's topology (cnn.c:416-428), or the conv -> ReLU -> max-pool variant (--pool).
inline std::vector<b200cnn_layer_desc> topology(bool pool) {
    if (!pool)
        return {{B200CNN_LAYER_INPUT, 1, 28, 28, 0, 0, 1}, {B200CNN_LAYER_CONV, 16, 14, 14, 3, 1, 2}, {B200CNN_LAYER_CONV, 32, 7, 7, 3, 1, 2},
                {B200CNN_LAYER_FULL, 200, 1, 1, 0, 0, 1},  {B200CNN_LAYER_FULL, 200, 1, 1, 0, 0, 1},  {B200CNN_LAYER_FULL, 10, 1, 1, 0, 0, 1}};
    return {{B200CNN_LAYER_INPUT, 1, 28, 28, 0, 0, 1}, {B200CNN_LAYER_CONV, 16, 28, 28, 3, 1, 1}, {B200CNN_LAYER_POOL, 16, 14, 14, 2, 0, 2},
            {B200CNN_LAYER_CONV, 32, 14, 14, 3, 1, 1}, {B200CNN_LAYER_POOL, 32, 7, 7, 2, 0, 2},   {B200CNN_LAYER_FULL, 200, 1, 1, 0, 0, 1},
            {B200CNN_LAYER_FULL, 200, 1, 1, 0, 0, 1},  {B200CNN_LAYER_FULL, 10, 1, 1, 0, 0, 1}};
}

// Initial parameters: the reference's generator and construction order (cnn.c:49, :323-340) through the CPU engine,
// or a checkpoint.  Returns 0 / 111.
inline int initial_params(const Options& o, size_t expect, std::vector<float>& p32) {
    GlibcRand rng(o.seed);
    Net cpu{Quirks{}};
    if (!o.pool) {
        cpu = Net::mnist(rng);
    } else {
        cpu.add_input(1, 28, 28);
        cpu.add_conv(16, 28, 28, 3, 1, 1, 0.1, rng);
        cpu.add_pool(2, 2);
        cpu.add_conv(32, 14, 14, 3, 1, 1, 0.1, rng);
        cpu.add_pool(2, 2);
        cpu.add_full(200, 0.1, rng);
        cpu.add_full(200, 0.1, rng);
        cpu.add_full(10, 0.1, rng);
    }
    std::vector<double> p64;
    cpu.gather_params(p64);
    p32.assign(p64.begin(), p64.end());
    if (p32.size() != expect) {
        fprintf(stderr, "internal error: CPU and GPU engines disagree on the parameter count (%zu vs %zu)\n", p32.size(), expect);
        return 111;
    }
    if (!o.load_path.empty()) {
        FILE* f = fopen(o.load_path.c_str(), "rb");
        const size_t got = f ? fread(p32.data(), sizeof(float), expect, f) : 0;
        const bool extra = f && fgetc(f) != EOF;
        if (f) fclose(f);
        if (!f || got != expect || extra) {
            fprintf(stderr, "cannot load %zu float32 parameters from %s\n", expect, o.load_path.c_str());
            return 111;
        }
    }
    return 0;
}

inline int save_params(const std::string& path, const std::vector<float>& p) {
    FILE* f = fopen(path.c_str(), "wb");
    const bool ok = f && fwrite(p.data(), sizeof(float), p.size(), f) == p.size();
    if (f) fclose(f);
    if (!ok) fprintf(stderr, "cannot write %s\n", path.c_str());
    return ok ? 0 : 111;
}

}  // namespace app
}  // namespace b200cnn
