// Mini-batch schedule of the CPU engine (see net.hpp).  The convolution / pooling prefix still runs one sample
// at a time through the single-sample kernels in net.cpp (their weights are a few KB and stay in L1); its
// activations are parked per sample so that backward can revisit them.  The fully connected tail is where the
// per-sample path spends its time waiting for memory, so it is restated as three register-blocked loops:
//   forward   Y[s][o]  = act(b[o] + sum_j W[o][j] X[s][j])            2 outputs x 4 samples per pass over j
//   backward  E[s][j]  = sum_o dnet[s][o] W[o][j]                     4 samples x 2 outputs per pass over j
//             dW[o][j] += sum_s dnet[s][o] X[s][j],  db[o] += sum_s   2 outputs x 4 samples per pass over j
// written with GCC vector extensions (4 doubles), compiled for AVX2 and baseline x86-64 like the other hot
// loops.  Replaces Layer_feedForw_full / Layer_feedBack_full (/root/reference/cnn.c:113-173) for whole batches.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "net.hpp"

namespace b200cnn {

#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__)
#define B200CNN_HOT __attribute__((target_clones("avx2", "default")))
#else
#define B200CNN_HOT
#endif

namespace {

typedef double v4d __attribute__((vector_size(32), aligned(8)));  // unaligned loads / stores allowed

// Macros, not functions: a function taking or returning a 32-byte vector would be compiled once for the baseline
// ISA (ABI warning, no AVX) instead of once per clone of its caller.
#define ld4(p) (*reinterpret_cast<const v4d*>(p))
#define st4(p, v) (*reinterpret_cast<v4d*>(p) = (v))
#define hsum(v) (((v)[0] + (v)[1]) + ((v)[2] + (v)[3]))
#define splat(x) (v4d{(x), (x), (x), (x)})


// ---- convolution as row-gather + register-blocked loops -----------------------------------------------------------
// rows[p][k]: for output pixel p the K = in_depth * k * k input values under the filter (zero where the filter hangs
// over the edge), k contiguous and padded to a multiple of 4.  With it the three convolution passes become
//   forward   y[oc][p]   = relu(b[oc] + rows[p] . W[oc])                  (dot products over k)
//   dW        dW[oc][:] += dnet[oc][p] * rows[p][:]                       (axpy over k, skipped where ReLU is off)
//   dX        drows[p][:] = sum_oc dnet[oc][p] * W[oc][:]   then scattered back to the input image
// all unit-stride over k whatever the convolution's stride, which is what the pixel-major loops in net.cpp
// (stride-2 reads, 7-pixel rows) cannot offer the vector units.
struct ConvGeom {
    int ic, ih, iw, oc, oh, ow, k, pad, stride, K, Kp, P;
};

ConvGeom conv_geom(const Layer& prev, const Layer& self) {
    ConvGeom g;
    g.ic = prev.depth; g.ih = prev.height; g.iw = prev.width;
    g.oc = self.depth; g.oh = self.height; g.ow = self.width;
    g.k = self.kernsize; g.pad = self.padding; g.stride = self.stride;
    g.K = g.ic * g.k * g.k;
    g.Kp = (g.K + 3) & ~3;
    g.P = g.oh * g.ow;
    return g;
}

// Gather table: tab[p * Kp + k] = offset of that filter tap's input value inside one CHW image, or `nin` (one
// past the end) where the tap hangs over the edge or k is row padding.  Images are staged with one extra zero
// element at [nin], so the gather and the scatter-add below are branch-free.
void build_gather_table(const ConvGeom& g, std::vector<int>& tab) {
    const int nin = g.ic * g.ih * g.iw;
    tab.assign(size_t(g.P) * g.Kp, nin);
    for (int oy = 0; oy < g.oh; oy++)
        for (int ox = 0; ox < g.ow; ox++) {
            int* r = &tab[size_t(oy * g.ow + ox) * g.Kp];
            int kidx = 0;
            for (int c = 0; c < g.ic; c++)
                for (int dy = 0; dy < g.k; dy++)
                    for (int dx = 0; dx < g.k; dx++, kidx++) {
                        const int y = oy * g.stride - g.pad + dy, x = ox * g.stride - g.pad + dx;
                        if (y >= 0 && y < g.ih && x >= 0 && x < g.iw) r[kidx] = (c * g.ih + y) * g.iw + x;
                    }
        }
}

// rows[i] = img[tab[i]]; img has nin + 1 elements, the last one zero.
void im2row(const std::vector<int>& tab, const double* img, double* rows) {
    const size_t n = tab.size();
    const int* t = tab.data();
    for (size_t i = 0; i < n; i++) rows[i] = img[t[i]];
}

// img[tab[i]] += drows[i]; contributions of padding land in the spare element.
void row2im_add(const std::vector<int>& tab, const double* drows, double* img) {
    const size_t n = tab.size();
    const int* t = tab.data();
    for (size_t i = 0; i < n; i++) img[t[i]] += drows[i];
}

// acc[0..NV) (vectors at dst) += sum over the listed i of coef[i * cstride] * src[i * sstride + 0..4NV).
// The list holds the i whose coefficient is not zero (about half are, behind a ReLU): compacting them first keeps
// the inner loop free of unpredictable branches.
template <int NV>
__attribute__((always_inline)) inline void axpy_rows(double* dst, const double* coef, size_t cstride, const double* src,
                                                     size_t sstride, const int* list, int count, bool load_dst) {
    v4d acc[NV];
    for (int v = 0; v < NV; v++) acc[v] = load_dst ? ld4(dst + 4 * v) : splat(0.0);
    for (int q = 0; q < count; q++) {
        const int i = list[q];
        const v4d k = splat(coef[size_t(i) * cstride]);
        const double* s = src + size_t(i) * sstride;
        for (int v = 0; v < NV; v++) acc[v] += k * ld4(s + 4 * v);
    }
    for (int v = 0; v < NV; v++) st4(dst + 4 * v, acc[v]);
}

__attribute__((always_inline)) inline void axpy_rows_n(int nv, double* dst, const double* coef, size_t cstride, const double* src,
                                                       size_t sstride, const int* list, int count, bool load_dst) {
    switch (nv) {
        case 8: axpy_rows<8>(dst, coef, cstride, src, sstride, list, count, load_dst); break;
        case 7: axpy_rows<7>(dst, coef, cstride, src, sstride, list, count, load_dst); break;
        case 6: axpy_rows<6>(dst, coef, cstride, src, sstride, list, count, load_dst); break;
        case 5: axpy_rows<5>(dst, coef, cstride, src, sstride, list, count, load_dst); break;
        case 4: axpy_rows<4>(dst, coef, cstride, src, sstride, list, count, load_dst); break;
        case 3: axpy_rows<3>(dst, coef, cstride, src, sstride, list, count, load_dst); break;
        case 2: axpy_rows<2>(dst, coef, cstride, src, sstride, list, count, load_dst); break;
        default: axpy_rows<1>(dst, coef, cstride, src, sstride, list, count, load_dst); break;
    }
}

// list = indices i in [0, count) with coef[i * stride] != 0, branch-free; returns how many.
inline int compact_nonzero(const double* coef, size_t stride, int count, int* list) {
    int m = 0;
    for (int i = 0; i < count; i++) {
        list[m] = i;
        m += coef[size_t(i) * stride] != 0.0;
    }
    return m;
}

B200CNN_HOT void conv_forward_rows(const ConvGeom& g, const double* rows, const double* wp, const double* bias, double* out,
                                   double* grad) {
    for (int p0 = 0; p0 < g.P; p0 += 2) {
        const int np = std::min(2, g.P - p0);
        const double* r0 = rows + size_t(p0) * g.Kp;
        const double* r1 = rows + size_t(p0 + np - 1) * g.Kp;
        for (int o0 = 0; o0 < g.oc; o0 += 4) {
            const int no = std::min(4, g.oc - o0);
            const double* w[4];
            for (int a = 0; a < 4; a++) w[a] = wp + size_t(o0 + std::min(a, no - 1)) * g.Kp;
            v4d a00 = splat(0.0), a01 = a00, a10 = a00, a11 = a00, a20 = a00, a21 = a00, a30 = a00, a31 = a00;
            for (int j = 0; j < g.Kp; j += 4) {
                const v4d x0 = ld4(r0 + j), x1 = ld4(r1 + j);
                const v4d u0 = ld4(w[0] + j), u1 = ld4(w[1] + j), u2 = ld4(w[2] + j), u3 = ld4(w[3] + j);
                a00 += u0 * x0; a01 += u0 * x1; a10 += u1 * x0; a11 += u1 * x1;
                a20 += u2 * x0; a21 += u2 * x1; a30 += u3 * x0; a31 += u3 * x1;
            }
            const double acc[4][2] = {{hsum(a00), hsum(a01)}, {hsum(a10), hsum(a11)}, {hsum(a20), hsum(a21)}, {hsum(a30), hsum(a31)}};
            for (int a = 0; a < no; a++)
                for (int b = 0; b < np; b++) {
                    const double v = bias[o0 + a] + acc[a][b];
                    const size_t idx = size_t(o0 + a) * g.P + p0 + b;
                    out[idx] = (0 < v) ? v : 0;
                    grad[idx] = (0 < v) ? 1 : 0;
                }
        }
    }
}

// dnet = err * grad; db += row sums of dnet; dwp += dnet x rows; drows = W^T dnet (when drows != NULL)
B200CNN_HOT void conv_backward_rows(const ConvGeom& g, const double* rows, const double* wp, const double* err, const double* grad,
                                    double* dnet, double* dwp, double* db, double* drows, int* list) {
    for (int o = 0; o < g.oc; o++) {
        double bsum = 0;
        for (int p = 0; p < g.P; p++) {
            const double d = err[size_t(o) * g.P + p] * grad[size_t(o) * g.P + p];
            dnet[size_t(o) * g.P + p] = d;
            bsum += d;
        }
        db[o] += bsum;
    }
    for (int o = 0; o < g.oc; o++) {
        const int m = compact_nonzero(dnet + size_t(o) * g.P, 1, g.P, list);
        for (int j0 = 0; j0 < g.Kp; j0 += 32)
            axpy_rows_n(std::min(8, (g.Kp - j0) / 4), dwp + size_t(o) * g.Kp + j0, dnet + size_t(o) * g.P, 1, rows + j0, size_t(g.Kp),
                        list, m, true);
    }
    if (drows)
        for (int p = 0; p < g.P; p++) {
            const int m = compact_nonzero(dnet + p, size_t(g.P), g.oc, list);
            for (int j0 = 0; j0 < g.Kp; j0 += 32)
                axpy_rows_n(std::min(8, (g.Kp - j0) / 4), drows + size_t(p) * g.Kp + j0, dnet + p, size_t(g.P), wp + j0, size_t(g.Kp),
                            list, m, false);
        }
}

}  // namespace

bool Net::batch_supported() const {
    return !(quirks_.shared_conv_slice || quirks_.early_update || quirks_.softmax_max_m1) && layers_.size() >= 2;
}

void Net::ensure_batch(int n) {
    if (n <= batch_cap_ && batch_.size() == layers_.size()) return;
    batch_cap_ = std::max(n, batch_cap_);
    batch_.resize(layers_.size());
    for (size_t li = 0; li < layers_.size(); li++) {
        const Layer& l = layers_[li];
        const size_t sz = size_t(batch_cap_) * l.nnodes();
        BatchBuf& b = batch_[li];
        b.outputs.assign(sz, 0.0);
        b.gradients.assign(sz, 0.0);
        b.errors.assign(sz, 0.0);
        if (l.type == LayerType::Pool) b.argmax.assign(sz, 0);
        if (l.type == LayerType::Conv) {
            const ConvGeom g = conv_geom(layers_[li - 1], l);
            b.wpad.assign(size_t(g.oc) * g.Kp, 0.0);
            b.dwpad.assign(size_t(g.oc) * g.Kp, 0.0);
            b.rows.assign(size_t(g.P) * g.Kp, 0.0);
            b.drows.assign(size_t(g.P) * g.Kp, 0.0);
            b.dnet.assign(size_t(g.oc) * g.P, 0.0);
            b.list.assign(size_t(std::max(g.oc, g.P)) + 1, 0);
            build_gather_table(g, b.gather);
            b.img.assign(size_t(layers_[li - 1].nnodes()) + 1, 0.0);
        }
    }
}

// Y[s][o] = act(b[o] + W[o] . X[s]) for the whole batch.  Two weight rows and four samples share one pass over j:
// 8 vector accumulators, 6 loads per 8 multiply-adds.
B200CNN_HOT void Net::forward_full_batch(size_t li, int n) {
    const Layer& self = layers_[li];
    const int nout = self.nnodes(), nin = layers_[li - 1].nnodes();
    const double* X = batch_[li - 1].outputs.data();
    double* Y = batch_[li].outputs.data();
    double* G = batch_[li].gradients.data();
    const double* W = self.weights.data();
    const int nin4 = nin & ~3;
    for (int o0 = 0; o0 < nout; o0 += 2) {
        const int no = std::min(2, nout - o0);
        const double* w0 = W + size_t(o0) * nin;
        const double* w1 = W + size_t(o0 + (no > 1 ? 1 : 0)) * nin;
        for (int s0 = 0; s0 < n; s0 += 4) {
            const int ns = std::min(4, n - s0);
            const double* x[4];
            for (int b = 0; b < 4; b++) x[b] = X + size_t(s0 + std::min(b, ns - 1)) * nin;
            v4d a00 = splat(0), a01 = a00, a02 = a00, a03 = a00, a10 = a00, a11 = a00, a12 = a00, a13 = a00;
            for (int j = 0; j < nin4; j += 4) {
                const v4d u0 = ld4(w0 + j), u1 = ld4(w1 + j);
                const v4d x0 = ld4(x[0] + j), x1 = ld4(x[1] + j), x2 = ld4(x[2] + j), x3 = ld4(x[3] + j);
                a00 += u0 * x0; a01 += u0 * x1; a02 += u0 * x2; a03 += u0 * x3;
                a10 += u1 * x0; a11 += u1 * x1; a12 += u1 * x2; a13 += u1 * x3;
            }
            double acc[2][4] = {{hsum(a00), hsum(a01), hsum(a02), hsum(a03)}, {hsum(a10), hsum(a11), hsum(a12), hsum(a13)}};
            for (int j = nin4; j < nin; j++)
                for (int b = 0; b < 4; b++) {
                    acc[0][b] += w0[j] * x[b][j];
                    acc[1][b] += w1[j] * x[b][j];
                }
            for (int a = 0; a < no; a++)
                for (int b = 0; b < ns; b++) Y[size_t(s0 + b) * nout + o0 + a] = self.biases[o0 + a] + acc[a][b];
        }
    }
    const bool last = li + 1 == layers_.size();
    for (int s = 0; s < n; s++) {
        double* y = Y + size_t(s) * nout;
        double* g = G + size_t(s) * nout;
        if (last) {  // softmax; its "gradient" is 1 (learn_* feeds softmax - onehot, cnn.c:125-143)
            double m = -std::numeric_limits<double>::infinity();
            for (int i = 0; i < nout; i++) m = std::max(m, y[i]);
            double t = 0;
            for (int i = 0; i < nout; i++) {
                y[i] = std::exp(y[i] - m);
                t += y[i];
            }
            for (int i = 0; i < nout; i++) {
                y[i] /= t;
                g[i] = 1;
            }
        } else {
            for (int i = 0; i < nout; i++) {
                const double v = std::tanh(y[i]);
                y[i] = v;
                g[i] = 1.0 - v * v;
            }
        }
    }
}

// errors of layer li (batch_[li].errors, [n][nout]) -> u_weights / u_biases of layer li and errors of layer li-1.
B200CNN_HOT void Net::backward_full_batch(size_t li, int n) {
    Layer& self = layers_[li];
    const int nout = self.nnodes(), nin = layers_[li - 1].nnodes();
    const double* X = batch_[li - 1].outputs.data();
    double* D = batch_[li].errors.data();  // becomes dnet = errors * gradients in place
    const double* G = batch_[li].gradients.data();
    for (size_t i = 0; i < size_t(n) * nout; i++) D[i] *= G[i];
    const double* W = self.weights.data();
    double* UW = self.u_weights.data();
    const int nin4 = nin & ~3;

    // data gradient (not needed below the first parameterised layer)
    if (li - 1 >= 1) {
        double* E = batch_[li - 1].errors.data();
        for (int s0 = 0; s0 < n; s0 += 4) {
            const int ns = std::min(4, n - s0);
            double* e[4];
            const double* d[4];
            for (int b = 0; b < 4; b++) {
                e[b] = E + size_t(s0 + std::min(b, ns - 1)) * nin;
                d[b] = D + size_t(s0 + std::min(b, ns - 1)) * nout;
            }
            for (int b = 0; b < ns; b++) std::memset(e[b], 0, sizeof(double) * nin);
            for (int o0 = 0; o0 < nout; o0 += 2) {
                const bool two = o0 + 1 < nout;
                const double* w0 = W + size_t(o0) * nin;
                const double* w1 = W + size_t(o0 + (two ? 1 : 0)) * nin;
                double c0[4], c1[4];
                for (int b = 0; b < 4; b++) {
                    c0[b] = d[b][o0];
                    c1[b] = two ? d[b][o0 + 1] : 0.0;
                }
                for (int b = 0; b < ns; b++) {  // (b >= ns aliases row ns-1: must not be accumulated twice)
                    const v4d k0 = splat(c0[b]), k1 = splat(c1[b]);
                    double* eb = e[b];
                    for (int j = 0; j < nin4; j += 4) st4(eb + j, ld4(eb + j) + k0 * ld4(w0 + j) + k1 * ld4(w1 + j));
                    for (int j = nin4; j < nin; j++) eb[j] += c0[b] * w0[j] + c1[b] * w1[j];
                }
            }
        }
    }

    // weight / bias gradient: two output rows x four samples per pass over j
    for (int o0 = 0; o0 < nout; o0 += 2) {
        const bool two = o0 + 1 < nout;
        double* u0 = UW + size_t(o0) * nin;
        double* u1 = UW + size_t(o0 + 1) * nin;  // only touched when `two`
        double b0 = 0, b1 = 0;
        for (int s0 = 0; s0 < n; s0 += 4) {
            const int ns = std::min(4, n - s0);
            const double* x[4];
            double c0[4], c1[4];
            for (int b = 0; b < 4; b++) {
                const int s = s0 + std::min(b, ns - 1);
                x[b] = X + size_t(s) * nin;
                c0[b] = b < ns ? D[size_t(s) * nout + o0] : 0.0;
                c1[b] = (b < ns && two) ? D[size_t(s) * nout + o0 + 1] : 0.0;
                b0 += c0[b];
                b1 += c1[b];
            }
            const v4d p0 = splat(c0[0]), p1 = splat(c0[1]), p2 = splat(c0[2]), p3 = splat(c0[3]);
            const v4d q0 = splat(c1[0]), q1 = splat(c1[1]), q2 = splat(c1[2]), q3 = splat(c1[3]);
            for (int j = 0; j < nin4; j += 4) {
                const v4d x0 = ld4(x[0] + j), x1 = ld4(x[1] + j), x2 = ld4(x[2] + j), x3 = ld4(x[3] + j);
                st4(u0 + j, ld4(u0 + j) + ((p0 * x0 + p1 * x1) + (p2 * x2 + p3 * x3)));
                if (two) st4(u1 + j, ld4(u1 + j) + ((q0 * x0 + q1 * x1) + (q2 * x2 + q3 * x3)));
            }
            for (int j = nin4; j < nin; j++)
                for (int b = 0; b < 4; b++) {
                    u0[j] += c0[b] * x[b][j];
                    if (two) u1[j] += c1[b] * x[b][j];
                }
        }
        self.u_biases[o0] += b0;
        if (two) self.u_biases[o0 + 1] += b1;
    }
}

void Net::forward_conv_batch(size_t li, int n) {
    const Layer& self = layers_[li];
    const ConvGeom g = conv_geom(layers_[li - 1], self);
    BatchBuf& b = batch_[li];
    for (int o = 0; o < g.oc; o++)  // refresh the padded copy of the weights (they change at every update)
        std::memcpy(&b.wpad[size_t(o) * g.Kp], &self.weights[size_t(o) * g.K], sizeof(double) * g.K);
    const size_t nin = size_t(layers_[li - 1].nnodes()), nout = size_t(self.nnodes());
    for (int s = 0; s < n; s++) {
        std::memcpy(b.img.data(), &batch_[li - 1].outputs[s * nin], sizeof(double) * nin);
        b.img[nin] = 0.0;  // the element every out-of-image tap reads (backward's scatter leaves its sum there)
        im2row(b.gather, b.img.data(), b.rows.data());
        conv_forward_rows(g, b.rows.data(), b.wpad.data(), self.biases.data(), &b.outputs[s * nout], &b.gradients[s * nout]);
    }
}

void Net::backward_conv_batch(size_t li, int n) {
    Layer& self = layers_[li];
    const ConvGeom g = conv_geom(layers_[li - 1], self);
    BatchBuf& b = batch_[li];
    const size_t nin = size_t(layers_[li - 1].nnodes()), nout = size_t(self.nnodes());
    const bool need_dx = li - 1 >= 1;  // nothing consumes the gradient w.r.t. the input image
    std::fill(b.dwpad.begin(), b.dwpad.end(), 0.0);
    for (int s = 0; s < n; s++) {
        std::memcpy(b.img.data(), &batch_[li - 1].outputs[s * nin], sizeof(double) * nin);
        b.img[nin] = 0.0;
        im2row(b.gather, b.img.data(), b.rows.data());
        conv_backward_rows(g, b.rows.data(), b.wpad.data(), &b.errors[s * nout], &b.gradients[s * nout], b.dnet.data(), b.dwpad.data(),
                           self.u_biases.data(), need_dx ? b.drows.data() : nullptr, b.list.data());
        if (need_dx) {
            std::fill(b.img.begin(), b.img.end(), 0.0);
            row2im_add(b.gather, b.drows.data(), b.img.data());
            std::memcpy(&batch_[li - 1].errors[s * nin], b.img.data(), sizeof(double) * nin);
        }
    }
    for (int o = 0; o < g.oc; o++)
        for (int k = 0; k < g.K; k++) self.u_weights[size_t(o) * g.K + k] += b.dwpad[size_t(o) * g.Kp + k];
}

void Net::forward_pool_batch(size_t li, int n) {
    const Layer& self = layers_[li];
    const Layer& prev = layers_[li - 1];
    const size_t nin = size_t(prev.nnodes()), nout = size_t(self.nnodes());
    for (int s = 0; s < n; s++) {
        const double* in = &batch_[li - 1].outputs[s * nin];
        double* out = &batch_[li].outputs[s * nout];
        double* grad = &batch_[li].gradients[s * nout];
        int* am = &batch_[li].argmax[s * nout];
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
                            if (best < 0 || bv < in[p]) {
                                best = p;
                                bv = in[p];
                            }
                        }
                    out[i] = bv;
                    grad[i] = 1;
                    am[i] = best;
                }
    }
}

void Net::backward_pool_batch(size_t li, int n) {
    const size_t nin = size_t(layers_[li - 1].nnodes()), nout = size_t(layers_[li].nnodes());
    for (int s = 0; s < n; s++) {
        double* din = &batch_[li - 1].errors[s * nin];
        std::memset(din, 0, sizeof(double) * nin);
        const double* e = &batch_[li].errors[s * nout];
        const int* am = &batch_[li].argmax[s * nout];
        for (size_t i = 0; i < nout; i++) din[am[i]] += e[i];
    }
}

void Net::forward_layers_batch(const double* xs, int n) {
    ensure_batch(n);
    std::memcpy(batch_[0].outputs.data(), xs, sizeof(double) * size_t(n) * layers_[0].nnodes());
    for (size_t li = 1; li < layers_.size(); li++) {
        switch (layers_[li].type) {
            case LayerType::Full: forward_full_batch(li, n); break;
            case LayerType::Conv: forward_conv_batch(li, n); break;
            case LayerType::Pool: forward_pool_batch(li, n); break;
            default: break;
        }
    }
}

void Net::learn_batch(const double* xs, const double* targets, int n, double* sample_error) {
    if (n <= 0) return;
    const size_t nin = size_t(layers_[0].nnodes()), nc = size_t(layers_.back().nnodes());
    if (!batch_supported()) {  // same result, one sample at a time
        for (int s = 0; s < n; s++) {
            set_inputs(xs + s * nin);
            learn_outputs(targets + s * nc);
            if (sample_error) sample_error[s] = error_total();
        }
        return;
    }
    const size_t L = layers_.size();
    forward_layers_batch(xs, n);
    // loss gradient w.r.t. the logits: softmax - onehot (cnn.c:285-286), and the reference's metric (cnn.c:275-282)
    {
        const double* Y = batch_[L - 1].outputs.data();
        double* E = batch_[L - 1].errors.data();
        for (int s = 0; s < n; s++) {
            double tot = 0;
            for (size_t i = 0; i < nc; i++) {
                const double e = Y[s * nc + i] - targets[s * nc + i];
                E[s * nc + i] = e;
                tot += e * e;
            }
            if (sample_error) sample_error[s] = tot / double(nc);
        }
    }
    for (size_t li = L - 1; li >= 1; li--) {
        switch (layers_[li].type) {
            case LayerType::Full: backward_full_batch(li, n); break;
            case LayerType::Conv: backward_conv_batch(li, n); break;
            case LayerType::Pool: backward_pool_batch(li, n); break;
            default: break;
        }
    }
    // leave the single-sample view consistent with the LAST sample (get_outputs / error_total callers)
    for (size_t li = 0; li < L; li++) {
        const size_t nn = size_t(layers_[li].nnodes());
        std::memcpy(layers_[li].outputs.data(), &batch_[li].outputs[(n - 1) * nn], sizeof(double) * nn);
    }
    for (size_t i = 0; i < nc; i++) layers_[L - 1].errors[i] = layers_[L - 1].outputs[i] - targets[(n - 1) * nc + i];
}

void Net::forward_batch(const double* xs, int n, double* ys) {
    if (n <= 0) return;
    const size_t nin = size_t(layers_[0].nnodes()), nc = size_t(layers_.back().nnodes());
    if (!batch_supported()) {
        for (int s = 0; s < n; s++) {
            set_inputs(xs + s * nin);
            get_outputs(ys + s * nc);
        }
        return;
    }
    forward_layers_batch(xs, n);
    std::memcpy(ys, batch_[layers_.size() - 1].outputs.data(), sizeof(double) * n * nc);
}

}  // namespace b200cnn
