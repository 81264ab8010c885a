#include "net.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>

namespace b200cnn {

// Hot loops are compiled twice (baseline x86-64 and AVX2) and dispatched at load time, so one binary runs on the
// build machine and on the GPU box.  -ffp-contract=off keeps both clones bit-identical (no FMA contraction).
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__)
#define B200CNN_HOT __attribute__((target_clones("avx2", "default")))
#else
#define B200CNN_HOT
#endif

Layer& Net::push(LayerType t, int d, int w, int h, size_t nb, size_t nw) {
    layers_.emplace_back();
    Layer& l = layers_.back();
    l.type = t;
    l.depth = d;
    l.width = w;
    l.height = h;
    const size_t n = size_t(d) * w * h;
    l.outputs.assign(n, 0.0);
    l.gradients.assign(n, 0.0);
    l.errors.assign(n, 0.0);
    l.biases.assign(nb, 0.0);
    l.u_biases.assign(nb, 0.0);
    l.weights.assign(nw, 0.0);
    l.u_weights.assign(nw, 0.0);
    return l;
}

int Net::add_input(int depth, int width, int height) {
    push(LayerType::Input, depth, width, height, 0, 0);
    return int(layers_.size()) - 1;
}

int Net::add_full(int nnodes, double std, GlibcRand& rng) {
    const size_t nin = size_t(layers_.back().nnodes());
    Layer& l = push(LayerType::Full, nnodes, 1, 1, size_t(nnodes), size_t(nnodes) * nin);
    for (double& w : l.weights) w = std * rng.nrnd();
    return int(layers_.size()) - 1;
}

int Net::add_conv(int depth, int width, int height, int kernsize, int padding, int stride, double std,
                  GlibcRand& rng) {
    const int in_depth = layers_.back().depth;
    Layer& l = push(LayerType::Conv, depth, width, height, size_t(depth),
                    size_t(depth) * in_depth * kernsize * kernsize);
    l.kernsize = kernsize;
    l.padding = padding;
    l.stride = stride;
    for (double& w : l.weights) w = std * rng.nrnd();
    return int(layers_.size()) - 1;
}

int Net::add_pool(int kernsize, int stride) {
    const Layer& p = layers_.back();
    const int ow = (p.width - kernsize) / stride + 1, oh = (p.height - kernsize) / stride + 1;
    const int d = p.depth;
    Layer& l = push(LayerType::Pool, d, ow, oh, 0, 0);
    l.kernsize = kernsize;
    l.stride = stride;
    l.argmax.assign(size_t(d) * ow * oh, 0);
    return int(layers_.size()) - 1;
}

Net Net::mnist(GlibcRand& rng, Quirks q) {
    Net net(q);
    net.add_input(1, 28, 28);
    net.add_conv(16, 14, 14, 3, 1, 2, 0.1, rng);
    net.add_conv(32, 7, 7, 3, 1, 2, 0.1, rng);
    net.add_full(200, 0.1, rng);
    net.add_full(200, 0.1, rng);
    net.add_full(10, 0.1, rng);
    return net;
}

// ---------------------------------------------------------------------------------------------
// Fully connected: y = act(W x + b), W row-major [out][in].  Hidden layers use tanh, the last
// layer of the net uses softmax whose "gradient" is 1 because learn_outputs feeds it
// (softmax - onehot), the exact cross-entropy gradient w.r.t. the logits.
// ---------------------------------------------------------------------------------------------
B200CNN_HOT void Net::forward_full(size_t li) {
    Layer& self = layers_[li];
    const Layer& prev = layers_[li - 1];
    const int nout = self.nnodes(), nin = prev.nnodes();
    const double* x = prev.outputs.data();
    const double* w = self.weights.data();
    const bool exact_order = quirks_.shared_conv_slice || quirks_.early_update || quirks_.softmax_max_m1;
    for (int i = 0; i < nout; i++, w += nin) {
        if (exact_order) {
            // One sequential accumulator: the summation order is part of the contract with the
            // reference oracle (bit-identical logits in quirks mode).
            double acc = self.biases[i];
            for (int j = 0; j < nin; j++) acc += x[j] * w[j];
            self.outputs[i] = acc;
        } else {
            // Eight independent partial sums: the dot product is add-latency bound otherwise.
            double a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0, a5 = 0, a6 = 0, a7 = 0;
            int j = 0;
            for (; j + 8 <= nin; j += 8) {
                a0 += x[j] * w[j];
                a1 += x[j + 1] * w[j + 1];
                a2 += x[j + 2] * w[j + 2];
                a3 += x[j + 3] * w[j + 3];
                a4 += x[j + 4] * w[j + 4];
                a5 += x[j + 5] * w[j + 5];
                a6 += x[j + 6] * w[j + 6];
                a7 += x[j + 7] * w[j + 7];
            }
            double acc = self.biases[i] + (((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7)));
            for (; j < nin; j++) acc += x[j] * w[j];
            self.outputs[i] = acc;
        }
    }
    if (li + 1 == layers_.size()) {
        double m = quirks_.softmax_max_m1 ? -1.0 : -std::numeric_limits<double>::infinity();
        for (int i = 0; i < nout; i++)
            if (m < self.outputs[i]) m = self.outputs[i];
        double t = 0;
        for (int i = 0; i < nout; i++) {
            const double y = std::exp(self.outputs[i] - m);
            self.outputs[i] = y;
            t += y;
        }
        for (int i = 0; i < nout; i++) {
            self.outputs[i] /= t;
            self.gradients[i] = 1;
        }
    } else {
        for (int i = 0; i < nout; i++) {
            const double y = std::tanh(self.outputs[i]);
            self.outputs[i] = y;
            self.gradients[i] = 1.0 - y * y;
        }
    }
}

B200CNN_HOT void Net::backward_full(size_t li) {
    Layer& self = layers_[li];
    Layer& prev = layers_[li - 1];
    const int nout = self.nnodes(), nin = prev.nnodes();
    double* perr = prev.errors.data();
    const double* pout = prev.outputs.data();
    for (int j = 0; j < nin; j++) perr[j] = 0;
    const double* w = self.weights.data();
    double* uw = self.u_weights.data();
    for (int i = 0; i < nout; i++, w += nin, uw += nin) {
        const double dnet = self.errors[i] * self.gradients[i];
        for (int j = 0; j < nin; j++) {
            perr[j] += w[j] * dnet;
            uw[j] += dnet * pout[j];
        }
        self.u_biases[i] += dnet;
    }
}

// ---------------------------------------------------------------------------------------------
// Convolution + ReLU, CHW single image, weights [out_c][in_c][ky][kx], zero padding.
// ---------------------------------------------------------------------------------------------
namespace {
// Valid output range [lo, hi) along one axis for filter offset d: 0 <= o*stride - pad + d < in.
inline void tap_range(int d, int stride, int pad, int in, int out, int& lo, int& hi) {
    lo = std::max(0, (pad - d + stride - 1) / stride);
    hi = std::min(out, (in - 1 + pad - d) / stride + 1);
}
}  // namespace

// Clean-mode convolution: one (out-channel, in-channel, tap) at a time over the whole output plane.  No bounds
// checks inside the pixel loops (the valid range per tap is computed once), so the compiler keeps them tight.
B200CNN_HOT void Net::forward_conv_fast(size_t li) {
    Layer& self = layers_[li];
    const Layer& prev = layers_[li - 1];
    const int k = self.kernsize, kk = k * k, s = self.stride, pad = self.padding;
    const int ow = self.width, oh = self.height, iw = prev.width, ih = prev.height;
    for (int z1 = 0; z1 < self.depth; z1++) {
        double* out = &self.outputs[size_t(z1) * ow * oh];
        std::fill(out, out + ow * oh, self.biases[z1]);
        for (int z0 = 0; z0 < prev.depth; z0++) {
            const double* in = &prev.outputs[size_t(z0) * iw * ih];
            const double* w = &self.weights[(size_t(z1) * prev.depth + z0) * kk];
            for (int dy = 0; dy < k; dy++) {
                int y_lo, y_hi;
                tap_range(dy, s, pad, ih, oh, y_lo, y_hi);
                for (int dx = 0; dx < k; dx++) {
                    int x_lo, x_hi;
                    tap_range(dx, s, pad, iw, ow, x_lo, x_hi);
                    const double wv = w[dy * k + dx];
                    for (int y1 = y_lo; y1 < y_hi; y1++) {
                        const double* irow = in + (y1 * s - pad + dy) * iw - pad + dx;
                        double* orow = out + y1 * ow;
                        for (int x1 = x_lo; x1 < x_hi; x1++) orow[x1] += irow[x1 * s] * wv;
                    }
                }
            }
        }
        double* g = &self.gradients[size_t(z1) * ow * oh];
        for (int i = 0; i < ow * oh; i++) {
            const double v = (0 < out[i]) ? out[i] : 0;
            out[i] = v;
            g[i] = (0 < v) ? 1 : 0;
        }
    }
}

B200CNN_HOT void Net::backward_conv_fast(size_t li) {
    Layer& self = layers_[li];
    Layer& prev = layers_[li - 1];
    std::fill(prev.errors.begin(), prev.errors.end(), 0.0);
    const int k = self.kernsize, kk = k * k, s = self.stride, pad = self.padding;
    const int ow = self.width, oh = self.height, iw = prev.width, ih = prev.height;
    std::vector<double>& dnet = scratch_;
    dnet.resize(size_t(ow) * oh);
    for (int z1 = 0; z1 < self.depth; z1++) {
        const double* err = &self.errors[size_t(z1) * ow * oh];
        const double* grad = &self.gradients[size_t(z1) * ow * oh];
        double bsum = 0;
        for (int i = 0; i < ow * oh; i++) {
            dnet[i] = err[i] * grad[i];
            bsum += dnet[i];
        }
        self.u_biases[z1] += bsum;
        for (int z0 = 0; z0 < prev.depth; z0++) {
            const double* in = &prev.outputs[size_t(z0) * iw * ih];
            double* perr = &prev.errors[size_t(z0) * iw * ih];
            const double* w = &self.weights[(size_t(z1) * prev.depth + z0) * kk];
            double* uw = &self.u_weights[(size_t(z1) * prev.depth + z0) * kk];
            for (int dy = 0; dy < k; dy++) {
                int y_lo, y_hi;
                tap_range(dy, s, pad, ih, oh, y_lo, y_hi);
                for (int dx = 0; dx < k; dx++) {
                    int x_lo, x_hi;
                    tap_range(dx, s, pad, iw, ow, x_lo, x_hi);
                    const double wv = w[dy * k + dx];
                    double acc = 0;
                    for (int y1 = y_lo; y1 < y_hi; y1++) {
                        const int off = (y1 * s - pad + dy) * iw - pad + dx;
                        const double* irow = in + off;
                        double* erow = perr + off;
                        const double* drow = dnet.data() + y1 * ow;
                        for (int x1 = x_lo; x1 < x_hi; x1++) {
                            erow[x1 * s] += wv * drow[x1];
                            acc += drow[x1] * irow[x1 * s];
                        }
                    }
                    uw[dy * k + dx] += acc;
                }
            }
        }
    }
}

void Net::forward_conv(size_t li) {
    if (!quirks_.shared_conv_slice) return forward_conv_fast(li);
    Layer& self = layers_[li];
    const Layer& prev = layers_[li - 1];
    const int k = self.kernsize, kk = k * k;
    const int zstride = quirks_.shared_conv_slice ? 0 : kk;  // D1 when 0
    int i = 0;
    for (int z1 = 0; z1 < self.depth; z1++) {
        const int qbase = z1 * prev.depth * kk;
        for (int y1 = 0; y1 < self.height; y1++) {
            const int y0 = self.stride * y1 - self.padding;
            for (int x1 = 0; x1 < self.width; x1++) {
                const int x0 = self.stride * x1 - self.padding;
                double v = self.biases[z1];
                for (int z0 = 0; z0 < prev.depth; z0++) {
                    const int pbase = z0 * prev.width * prev.height;
                    for (int dy = 0; dy < k; dy++) {
                        const int y = y0 + dy;
                        if (y < 0 || y >= prev.height) continue;
                        const int p = pbase + y * prev.width;
                        const int q = qbase + z0 * zstride + dy * k;
                        for (int dx = 0; dx < k; dx++) {
                            const int x = x0 + dx;
                            if (0 <= x && x < prev.width) v += prev.outputs[p + x] * self.weights[q + dx];
                        }
                    }
                }
                v = (0 < v) ? v : 0;
                self.outputs[i] = v;
                self.gradients[i] = (0 < v) ? 1 : 0;
                i++;
            }
        }
    }
}

void Net::backward_conv(size_t li) {
    if (!quirks_.shared_conv_slice) return backward_conv_fast(li);
    Layer& self = layers_[li];
    Layer& prev = layers_[li - 1];
    for (double& e : prev.errors) e = 0;
    const int k = self.kernsize, kk = k * k;
    const int zstride = quirks_.shared_conv_slice ? 0 : kk;
    int i = 0;
    for (int z1 = 0; z1 < self.depth; z1++) {
        const int qbase = z1 * prev.depth * kk;
        for (int y1 = 0; y1 < self.height; y1++) {
            const int y0 = self.stride * y1 - self.padding;
            for (int x1 = 0; x1 < self.width; x1++) {
                const int x0 = self.stride * x1 - self.padding;
                const double dnet = self.errors[i] * self.gradients[i];
                for (int z0 = 0; z0 < prev.depth; z0++) {
                    const int pbase = z0 * prev.width * prev.height;
                    for (int dy = 0; dy < k; dy++) {
                        const int y = y0 + dy;
                        if (y < 0 || y >= prev.height) continue;
                        const int p = pbase + y * prev.width;
                        const int q = qbase + z0 * zstride + dy * k;
                        for (int dx = 0; dx < k; dx++) {
                            const int x = x0 + dx;
                            if (0 <= x && x < prev.width) {
                                prev.errors[p + x] += self.weights[q + dx] * dnet;
                                self.u_weights[q + dx] += dnet * prev.outputs[p + x];
                            }
                        }
                    }
                }
                self.u_biases[z1] += dnet;
                i++;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Max pooling (no parameters).  Not part of the reference's network; provided because the
// project's stated topology is conv -> activation -> pool -> FC -> softmax.
// ---------------------------------------------------------------------------------------------
void Net::forward_pool(size_t li) {
    Layer& self = layers_[li];
    const Layer& prev = layers_[li - 1];
    int i = 0;
    for (int z = 0; z < self.depth; z++)
        for (int y1 = 0; y1 < self.height; y1++)
            for (int x1 = 0; x1 < self.width; x1++, i++) {
                int best = -1;
                double bv = 0;
                for (int dy = 0; dy < self.kernsize; dy++)
                    for (int dx = 0; dx < self.kernsize; dx++) {
                        const int y = y1 * self.stride + dy, x = x1 * self.stride + dx;
                        const int p = (z * prev.height + y) * prev.width + x;
                        if (best < 0 || bv < prev.outputs[p]) {
                            best = p;
                            bv = prev.outputs[p];
                        }
                    }
                self.outputs[i] = bv;
                self.gradients[i] = 1;
                self.argmax[i] = best;
            }
}

void Net::backward_pool(size_t li) {
    Layer& self = layers_[li];
    Layer& prev = layers_[li - 1];
    for (double& e : prev.errors) e = 0;
    for (int i = 0; i < self.nnodes(); i++) prev.errors[self.argmax[i]] += self.errors[i];
}

void Net::set_inputs(const double* x) {
    Layer& in = layers_[0];
    std::memcpy(in.outputs.data(), x, sizeof(double) * in.nnodes());
    for (size_t li = 1; li < layers_.size(); li++) {
        switch (layers_[li].type) {
            case LayerType::Full: forward_full(li); break;
            case LayerType::Conv: forward_conv(li); break;
            case LayerType::Pool: forward_pool(li); break;
            default: break;
        }
    }
}

void Net::get_outputs(double* y) const {
    const Layer& out = layers_.back();
    std::memcpy(y, out.outputs.data(), sizeof(double) * out.nnodes());
}

void Net::learn_outputs(const double* target) {
    Layer& out = layers_.back();
    for (int i = 0; i < out.nnodes(); i++) out.errors[i] = out.outputs[i] - target[i];
    for (size_t li = layers_.size() - 1; li >= 1; li--) {
        switch (layers_[li].type) {
            case LayerType::Full: backward_full(li); break;
            case LayerType::Conv: backward_conv(li); break;
            case LayerType::Pool: backward_pool(li); break;
            default: break;
        }
    }
}

double Net::error_total() const {
    const Layer& out = layers_.back();
    double total = 0;
    for (double e : out.errors) total += e * e;
    return total / out.nnodes();
}

void Net::update(double rate) {
    // Output layer first, like the reference's recursion; the order is irrelevant numerically.
    for (size_t li = layers_.size(); li-- > 0;) {
        Layer& l = layers_[li];
        for (size_t i = 0; i < l.biases.size(); i++) {
            l.biases[i] -= rate * l.u_biases[i];
            l.u_biases[i] = 0;
        }
        for (size_t i = 0; i < l.weights.size(); i++) {
            l.weights[i] -= rate * l.u_weights[i];
            l.u_weights[i] = 0;
        }
    }
}

void Net::zero_updates() {
    for (Layer& l : layers_) {
        std::fill(l.u_biases.begin(), l.u_biases.end(), 0.0);
        std::fill(l.u_weights.begin(), l.u_weights.end(), 0.0);
    }
}

size_t Net::num_params() const {
    size_t n = 0;
    for (const Layer& l : layers_) n += l.weights.size() + l.biases.size();
    return n;
}

void Net::gather_updates(std::vector<double>& flat) const {
    flat.clear();
    flat.reserve(num_params());
    for (const Layer& l : layers_) {
        flat.insert(flat.end(), l.u_weights.begin(), l.u_weights.end());
        flat.insert(flat.end(), l.u_biases.begin(), l.u_biases.end());
    }
}

void Net::scatter_updates(const std::vector<double>& flat) {
    size_t o = 0;
    for (Layer& l : layers_) {
        std::copy(flat.begin() + o, flat.begin() + o + l.u_weights.size(), l.u_weights.begin());
        o += l.u_weights.size();
        std::copy(flat.begin() + o, flat.begin() + o + l.u_biases.size(), l.u_biases.begin());
        o += l.u_biases.size();
    }
}

void Net::gather_params(std::vector<double>& flat) const {
    flat.clear();
    flat.reserve(num_params());
    for (const Layer& l : layers_) {
        flat.insert(flat.end(), l.weights.begin(), l.weights.end());
        flat.insert(flat.end(), l.biases.begin(), l.biases.end());
    }
}

void Net::scatter_params(const std::vector<double>& flat) {
    size_t o = 0;
    for (Layer& l : layers_) {
        std::copy(flat.begin() + o, flat.begin() + o + l.weights.size(), l.weights.begin());
        o += l.weights.size();
        std::copy(flat.begin() + o, flat.begin() + o + l.biases.size(), l.biases.begin());
        o += l.biases.size();
    }
}

// Checkpoint: "B2CN" magic, version, parameter count, then fp64 params in gather_params order.
bool Net::save(const std::string& path, std::string* err) const {
    FILE* fp = fopen(path.c_str(), "wb");
    if (!fp) {
        if (err) *err = "cannot create " + path;
        return false;
    }
    std::vector<double> flat;
    gather_params(flat);
    const uint32_t head[2] = {0x4e433242u, 1u};
    const uint64_t n = flat.size();
    bool ok = fwrite(head, sizeof(head), 1, fp) == 1 && fwrite(&n, sizeof(n), 1, fp) == 1 &&
              fwrite(flat.data(), sizeof(double), flat.size(), fp) == flat.size();
    fclose(fp);
    if (!ok && err) *err = "write failed: " + path;
    return ok;
}

bool Net::load(const std::string& path, std::string* err) {
    FILE* fp = fopen(path.c_str(), "rb");
    if (!fp) {
        if (err) *err = "cannot open " + path;
        return false;
    }
    uint32_t head[2] = {0, 0};
    uint64_t n = 0;
    std::vector<double> flat;
    bool ok = fread(head, sizeof(head), 1, fp) == 1 && fread(&n, sizeof(n), 1, fp) == 1 &&
              head[0] == 0x4e433242u && head[1] == 1u && n == num_params();
    if (ok) {
        flat.resize(n);
        ok = fread(flat.data(), sizeof(double), n, fp) == n;
    }
    fclose(fp);
    if (!ok) {
        if (err) *err = "not a matching checkpoint: " + path;
        return false;
    }
    scatter_params(flat);
    return true;
}

}  // namespace b200cnn
