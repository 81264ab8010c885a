// extern "C" surface of the CPU engine; see include/b200cnn_cpu.h.
#include <cstring>

#include "../../include/b200cnn_cpu.h"
#include "idx.hpp"
#include "net.hpp"

using namespace b200cnn;

struct b200cnn_cpu_net {
    Net net;
    GlibcRand rng;
    explicit b200cnn_cpu_net(Quirks q) : net(q), rng(0) {}
};

extern "C" {

b200cnn_cpu_net* b200cnn_cpu_net_create(unsigned quirks) {
    Quirks q;
    q.shared_conv_slice = quirks & 1u;
    q.early_update = quirks & 2u;
    q.softmax_max_m1 = quirks & 4u;
    return new b200cnn_cpu_net(q);
}
void b200cnn_cpu_net_destroy(b200cnn_cpu_net* n) { delete n; }
void b200cnn_cpu_net_seed(b200cnn_cpu_net* n, unsigned seed) { n->rng.reseed(seed); }

int b200cnn_cpu_net_add_input(b200cnn_cpu_net* n, int d, int w, int h) { return n->net.add_input(d, w, h); }
int b200cnn_cpu_net_add_conv(b200cnn_cpu_net* n, int d, int w, int h, int k, int pad, int stride, double std) {
    return n->net.add_conv(d, w, h, k, pad, stride, std, n->rng);
}
int b200cnn_cpu_net_add_full(b200cnn_cpu_net* n, int nnodes, double std) {
    return n->net.add_full(nnodes, std, n->rng);
}
int b200cnn_cpu_net_add_pool(b200cnn_cpu_net* n, int k, int stride) { return n->net.add_pool(k, stride); }

int b200cnn_cpu_net_num_layers(const b200cnn_cpu_net* n) { return int(n->net.num_layers()); }
int b200cnn_cpu_net_layer_info(const b200cnn_cpu_net* n, int li, int* info) {
    const Layer& l = n->net.layer(size_t(li));
    info[0] = int(l.type);
    info[1] = l.depth;
    info[2] = l.width;
    info[3] = l.height;
    info[4] = l.kernsize;
    info[5] = l.padding;
    info[6] = l.stride;
    info[7] = int(l.weights.size());
    return int(l.biases.size());
}
size_t b200cnn_cpu_net_num_params(const b200cnn_cpu_net* n) { return n->net.num_params(); }

void b200cnn_cpu_net_set_inputs(b200cnn_cpu_net* n, const double* x) { n->net.set_inputs(x); }
void b200cnn_cpu_net_get_outputs(const b200cnn_cpu_net* n, double* y) { n->net.get_outputs(y); }
void b200cnn_cpu_net_learn_outputs(b200cnn_cpu_net* n, const double* t) { n->net.learn_outputs(t); }
double b200cnn_cpu_net_error_total(const b200cnn_cpu_net* n) { return n->net.error_total(); }
void b200cnn_cpu_net_update(b200cnn_cpu_net* n, double rate) { n->net.update(rate); }
int b200cnn_cpu_net_batch_supported(const b200cnn_cpu_net* n) { return n->net.batch_supported() ? 1 : 0; }
void b200cnn_cpu_net_learn_batch(b200cnn_cpu_net* n, const double* xs, const double* targets, int count, double* sample_error) {
    n->net.learn_batch(xs, targets, count, sample_error);
}
void b200cnn_cpu_net_forward_batch(b200cnn_cpu_net* n, const double* xs, int count, double* ys) {
    n->net.forward_batch(xs, count, ys);
}

void b200cnn_cpu_net_get_params(const b200cnn_cpu_net* n, double* out) {
    std::vector<double> flat;
    n->net.gather_params(flat);
    std::memcpy(out, flat.data(), flat.size() * sizeof(double));
}
void b200cnn_cpu_net_set_params(b200cnn_cpu_net* n, const double* in) {
    std::vector<double> flat(in, in + n->net.num_params());
    n->net.scatter_params(flat);
}
void b200cnn_cpu_net_get_updates(const b200cnn_cpu_net* n, double* out) {
    std::vector<double> flat;
    n->net.gather_updates(flat);
    std::memcpy(out, flat.data(), flat.size() * sizeof(double));
}
int b200cnn_cpu_net_layer_array(const b200cnn_cpu_net* n, int li, int which, double* out) {
    const Layer& l = n->net.layer(size_t(li));
    const std::vector<double>& v = which == 0 ? l.outputs : (which == 1 ? l.gradients : l.errors);
    std::memcpy(out, v.data(), v.size() * sizeof(double));
    return int(v.size());
}

int b200cnn_cpu_net_save(const b200cnn_cpu_net* n, const char* path) { return n->net.save(path) ? 0 : 111; }
int b200cnn_cpu_net_load(b200cnn_cpu_net* n, const char* path) { return n->net.load(path) ? 0 : 111; }

void b200cnn_glibc_rand(unsigned seed, int n, int* out) {
    GlibcRand r(seed);
    for (int i = 0; i < n; i++) out[i] = r.next();
}
int b200cnn_cpu_net_rand(b200cnn_cpu_net* n) { return n->rng.next(); }

void b200cnn_make_synthetic(unsigned n, unsigned h, unsigned w, unsigned nclass, uint64_t seed, uint8_t* images,
                            uint8_t* labels) {
    IdxFile x, y;
    make_synthetic(n, h, w, nclass, seed, x, y);
    std::memcpy(images, x.data.data(), x.data.size());
    std::memcpy(labels, y.data.data(), y.data.size());
}

int b200cnn_idx_write(const char* path, int ndims, const uint32_t* dims, const uint8_t* data) {
    IdxFile f;
    f.dims.assign(dims, dims + ndims);
    size_t n = 1;
    for (uint32_t d : f.dims) n *= d;
    f.data.assign(data, data + n);
    return idx_write(path, f) ? 0 : 111;
}

int b200cnn_idx_read(const char* path, int* ndims, uint32_t* dims, uint8_t* data, size_t capacity) {
    IdxFile f;
    if (!idx_read(path, f)) return 111;
    *ndims = int(f.dims.size());
    for (size_t i = 0; i < f.dims.size() && i < 4; i++) dims[i] = f.dims[i];
    if (data) {
        if (capacity < f.data.size()) return 111;
        std::memcpy(data, f.data.data(), f.data.size());
    }
    return 0;
}

}  // extern "C"
