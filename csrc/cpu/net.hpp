// CPU network engine (fp64): the native counterpart of the reference's Layer API.
//
// Reference surface being re-provided (all in /root/reference/cnn.c):
//   struct Layer + Layer_create/destroy            :15-43, :60-107      -> b200cnn::Layer / Net
//   Layer_create_input / _full / _conv             :316-342             -> Net::add_input/add_full/add_conv
//   Layer_feedForw_full / feedBack_full            :113-173             -> Net::forward_full / backward_full
//   Layer_feedForw_conv / feedBack_conv            :175-247             -> Net::forward_conv / backward_conv
//   Layer_setInputs / getOutputs / learnOutputs    :249-301             -> Net::set_inputs / get_outputs / learn_outputs
//   Layer_getErrorTotal / Layer_update             :275-282, :303-314   -> Net::error_total / update
//
// Design differences (deliberate): layers live in one vector owned by a Net instead of a
// hand-linked list of calloc'd nodes; a max-pool layer type exists (the north star's topology
// names one, the reference's hard-coded network does not use one); the conv indexes the full
// [out_c][in_c][ky][kx] weight tensor.  `Quirks` re-enables the reference's defects D1/D2/D4
//  so this engine can be diffed bit-for-bit against the unmodified cnn.c.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "glibc_rand.hpp"

namespace b200cnn {

enum class LayerType : int { Input = 0, Full = 1, Conv = 2, Pool = 3 };

struct Quirks {
    bool shared_conv_slice = false;  // D1: conv weight index ignores the input channel
    bool early_update = false;       // D2: SGD fires when i % batch == 0 (first batch = 1 sample, tail dropped)
    bool softmax_max_m1 = false;     // D4: running max of the softmax starts at -1 instead of -inf
    static Quirks reference() { return Quirks{true, true, true}; }
};

struct Layer {
    LayerType type = LayerType::Input;
    int depth = 0, width = 0, height = 0;
    int kernsize = 0, padding = 0, stride = 1;  // conv / pool only
    std::vector<double> outputs;    // activations, CHW
    std::vector<double> gradients;  // d activation / d pre-activation (the reference's name)
    std::vector<double> errors;     // dLoss / d outputs
    std::vector<double> biases, u_biases;
    std::vector<double> weights, u_weights;
    std::vector<int> argmax;        // pool only: winning input index per output
    int nnodes() const { return depth * width * height; }
};

class Net {
public:
    explicit Net(Quirks q = Quirks{}) : quirks_(q) {}

    // Constructors.  Weights are std * nrnd() drawn from `rng` in layer order, 4 draws per weight
    // (cnn.c:49, :323-324, :339-340); biases start at 0.
    int add_input(int depth, int width, int height);
    int add_full(int nnodes, double std, GlibcRand& rng);
    int add_conv(int depth, int width, int height, int kernsize, int padding, int stride, double std,
                 GlibcRand& rng);
    int add_pool(int kernsize, int stride);  // max pool, output size floor((in - k) / stride) + 1

    // The reference's hard-coded topology (cnn.c:416-428).
    static Net mnist(GlibcRand& rng, Quirks q = Quirks{});

    // Forward a single CHW sample through every layer (cnn.c:249-268).
    void set_inputs(const double* x);
    void get_outputs(double* y) const;  // last layer, cnn.c:270-273
    // errors = outputs - target on the last layer, then backward through all layers; parameter
    // gradients ACCUMULATE into u_weights / u_biases (cnn.c:284-301).
    void learn_outputs(const double* target);
    double error_total() const;  // mean squared error of the last layer's errors (cnn.c:275-282)
    void update(double rate);    // w -= rate * u_w; u_w = 0 for every layer (cnn.c:303-314)
    void zero_updates();

    size_t num_layers() const { return layers_.size(); }
    Layer& layer(size_t i) { return layers_[i]; }
    const Layer& layer(size_t i) const { return layers_[i]; }
    size_t num_params() const;
    // Flat parameter / gradient views in layer order, weights then biases per layer — the order
    // the data-parallel trainer all-reduces and the checkpoint file stores.
    void gather_updates(std::vector<double>& flat) const;
    void scatter_updates(const std::vector<double>& flat);
    void gather_params(std::vector<double>& flat) const;
    void scatter_params(const std::vector<double>& flat);

    bool save(const std::string& path, std::string* err = nullptr) const;
    bool load(const std::string& path, std::string* err = nullptr);

    const Quirks& quirks() const { return quirks_; }

    // ---- mini-batch fast path (net_batch.cpp) ------------------------------------------------------------
    // learn_batch(xs, targets, n) is n calls of set_inputs + learn_outputs: gradients ACCUMULATE into
    // u_weights / u_biases, sample_error[s] is what error_total() would have returned after sample s.  The
    // difference is the schedule: the fully connected layers run as matrix-matrix loops over the whole batch, so
    // their weights (2.5 MB for the first one: larger than L2) are streamed once per batch instead of three times
    // per sample, which is what bounds the per-sample path.  Only the summation ORDER differs (results agree to
    // rounding); convolutions run per sample as a row gather + register-blocked unit-stride loops.  Clean mode only.
    bool batch_supported() const;
    void learn_batch(const double* xs, const double* targets, int n, double* sample_error = nullptr);
    void forward_batch(const double* xs, int n, double* ys);  // ys: [n][outputs of the last layer]

private:
    void forward_full(size_t li);
    void backward_full(size_t li);
    void forward_conv(size_t li);
    void backward_conv(size_t li);
    void forward_conv_fast(size_t li);   // clean mode: tap-outermost, branch-free pixel loops
    void backward_conv_fast(size_t li);
    void forward_pool(size_t li);
    void backward_pool(size_t li);
    Layer& push(LayerType t, int d, int w, int h, size_t nb, size_t nw);

    std::vector<Layer> layers_;
    Quirks quirks_;
    std::vector<double> scratch_;  // dLoss/d(pre-activation) of one output plane (backward_conv_fast)

    struct BatchBuf {
        std::vector<double> outputs, gradients, errors;  // [capacity][nnodes]
        std::vector<int> argmax;                          // pool layers: [capacity][nnodes]
        // convolution layers: weights / weight-gradient accumulators with rows padded to a multiple of 4, and the
        // per-sample row-gather scratch (see net_batch.cpp)
        std::vector<double> wpad, dwpad, rows, drows, dnet, img;
        std::vector<int> gather, list;
    };
    std::vector<BatchBuf> batch_;  // one per layer
    int batch_cap_ = 0;
    void ensure_batch(int n);
    void forward_layers_batch(const double* xs, int n);
    void forward_full_batch(size_t li, int n);
    void backward_full_batch(size_t li, int n);
    void forward_conv_batch(size_t li, int n);
    void backward_conv_batch(size_t li, int n);
    void forward_pool_batch(size_t li, int n);
    void backward_pool_batch(size_t li, int n);
};

}  // namespace b200cnn
