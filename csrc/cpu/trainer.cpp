#include "trainer.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace b200cnn {

namespace {

void usage(const char* prog, bool dp) {
    fprintf(stderr,
            "usage: %s [options] train-images train-labels test-images test-labels\n"
            "  --rate R         learning rate (default 0.1)\n"
            "  --nepoch N       passes over the training set (default 10)\n"
            "  --batch-size B   samples per SGD update%s (default 32)\n"
            "  --seed S         seed of the glibc-compatible generator (default 0)\n"
            "  --log-every N    samples between progress lines (default 1000)\n"
            "  --synthetic N    train on N generated MNIST-shaped samples; the paths become optional\n"
            "  --synthetic-test N\n"
            "  --pool           conv(3x3,s1)+ReLU+maxpool(2x2) blocks instead of stride-2 convs\n"
            "  --ref-quirks     reproduce the reference's conv indexing / batching quirks\n"
            "  --save F / --load F   checkpoint the parameters\n"
            "  --per-sample          clean mode: one sample at a time through the layers (default: mini-batch schedule)\n"
            "  --timing         report training images/s\n%s",
            prog, dp ? " per rank" : "", dp ? "  -np N            fork N ranks over shared memory\n" : "");
}

double now_s() {
    using clk = std::chrono::steady_clock;
    return std::chrono::duration<double>(clk::now().time_since_epoch()).count();
}

void load_sample(const IdxFile& images, const IdxFile& labels, size_t index, double* x, double* onehot,
                 int nclass) {
    const uint8_t* img = images.get3(index);
    const size_t n = images.item_size();
    for (size_t j = 0; j < n; j++) x[j] = img[j] / 255.0;  // cnn.c:457
    const int label = labels.get1(index);
    for (int j = 0; j < nclass; j++) onehot[j] = (j == label) ? 1 : 0;
}

}  // namespace

int parse_options(int argc, char** argv, bool dp, TrainOptions& opt) {
    std::vector<std::string> pos;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto need = [&](const char* name) -> const char* {
            if (i + 1 >= argc) {
                fprintf(stderr, "%s needs a value\n", name);
                return nullptr;
            }
            return argv[++i];
        };
        const char* v = nullptr;
        if (a == "--rate") { if (!(v = need("--rate"))) return kUsage; opt.rate = atof(v); }
        else if (a == "--nepoch") { if (!(v = need("--nepoch"))) return kUsage; opt.nepoch = atoi(v); }
        else if (a == "--batch-size") { if (!(v = need("--batch-size"))) return kUsage; opt.batch_size = atoi(v); }
        else if (a == "--seed") { if (!(v = need("--seed"))) return kUsage; opt.seed = unsigned(strtoul(v, nullptr, 10)); }
        else if (a == "--log-every") { if (!(v = need("--log-every"))) return kUsage; opt.log_every = atoi(v); }
        else if (a == "--synthetic") { if (!(v = need("--synthetic"))) return kUsage; opt.synthetic = atoi(v); }
        else if (a == "--synthetic-test") { if (!(v = need("--synthetic-test"))) return kUsage; opt.synthetic_test = atoi(v); }
        else if (a == "--save") { if (!(v = need("--save"))) return kUsage; opt.save_path = v; }
        else if (a == "--load") { if (!(v = need("--load"))) return kUsage; opt.load_path = v; }
        else if (a == "--ref-quirks") opt.ref_quirks = true;
        else if (a == "--per-sample") opt.per_sample = true;
        else if (a == "--pool") opt.pool = true;
        else if (a == "--timing") opt.timing = true;
        else if (dp && (a == "-np" || a == "--np")) { if (!(v = need("-np"))) return kUsage; opt.np = atoi(v); }
        else if (a == "-h" || a == "--help") { usage(argv[0], dp); return kUsage; }
        else if (a.size() > 1 && a[0] == '-' && a[1] == '-') { fprintf(stderr, "unknown option %s\n", a.c_str()); usage(argv[0], dp); return kUsage; }
        else pos.push_back(a);
    }
    if (opt.batch_size < 1 || opt.nepoch < 0 || opt.log_every < 1) return kUsage;
    if (opt.synthetic <= 0 && pos.size() < 4) return kUsage;  // the reference tests argc<4 (D3)
    if (pos.size() >= 4) {
        opt.train_images = pos[0];
        opt.train_labels = pos[1];
        opt.test_images = pos[2];
        opt.test_labels = pos[3];
    }
    return kOk;
}

int load_data(const TrainOptions& opt, DataSet& data) {
    if (opt.synthetic > 0 && opt.train_images.empty()) {
        int ntest = opt.synthetic_test > 0 ? opt.synthetic_test : std::max(100, opt.synthetic / 5);
        make_synthetic(uint32_t(opt.synthetic), 28, 28, 10, 1, data.train_x, data.train_y);
        make_synthetic(uint32_t(ntest), 28, 28, 10, 2, data.test_x, data.test_y);
        return kOk;
    }
    std::string err;
    const std::string* paths[4] = {&opt.train_images, &opt.train_labels, &opt.test_images, &opt.test_labels};
    IdxFile* files[4] = {&data.train_x, &data.train_y, &data.test_x, &data.test_y};
    for (int i = 0; i < 4; i++)
        if (!idx_read(*paths[i], *files[i], &err)) {
            fprintf(stderr, "%s\n", err.c_str());
            return kFileError;
        }
    if (data.train_x.dims.size() != 3 || data.train_x.count() != data.train_y.count() ||
        data.test_x.dims.size() != 3 || data.test_x.count() != data.test_y.count()) {
        fprintf(stderr, "image / label files do not match\n");
        return kFileError;
    }
    // The network reads item_size() bytes per image: a smaller or larger image file would make set_inputs read past
    // the sample (the reference memcpy's a fixed 28*28, cnn.c:399-402, and checks nothing).
    if (data.train_y.dims.size() != 1 || data.test_y.dims.size() != 1 || data.train_x.item_size() != 28 * 28 ||
        data.test_x.item_size() != 28 * 28) {
        fprintf(stderr, "expected 28x28 images (idx3) and 1-D label files (idx1)\n");
        return kFileError;
    }
    return kOk;
}

Net build_net(const TrainOptions& opt, GlibcRand& rng) {
    const Quirks q = opt.ref_quirks ? Quirks::reference() : Quirks{};
    if (!opt.pool) return Net::mnist(rng, q);
    Net net(q);
    net.add_input(1, 28, 28);
    net.add_conv(16, 28, 28, 3, 1, 1, 0.1, rng);
    net.add_pool(2, 2);
    net.add_conv(32, 14, 14, 3, 1, 1, 0.1, rng);
    net.add_pool(2, 2);
    net.add_full(200, 0.1, rng);
    net.add_full(200, 0.1, rng);
    net.add_full(10, 0.1, rng);
    return net;
}

int evaluate(Net& net, const IdxFile& images, const IdxFile& labels, bool log) {
    const int ntests = int(images.count());
    const int nclass = net.layer(net.num_layers() - 1).nnodes();
    std::vector<double> x(images.item_size()), y(nclass);
    int ncorrect = 0;
    if (net.batch_supported()) {
        // clean mode: 64 samples per pass over the fully connected weights
        const int chunk = 64;
        const size_t item = images.item_size();
        std::vector<double> xs(size_t(chunk) * item), ys(size_t(chunk) * nclass);
        for (int i0 = 0; i0 < ntests; i0 += chunk) {
            const int nb = std::min(chunk, ntests - i0);
            for (int b = 0; b < nb; b++) {
                const uint8_t* img = images.get3(i0 + b);
                for (size_t j = 0; j < item; j++) xs[b * item + j] = img[j] / 255.0;
            }
            net.forward_batch(xs.data(), nb, ys.data());
            for (int b = 0; b < nb; b++) {
                const double* yb = &ys[size_t(b) * nclass];
                int mj = -1;
                for (int j = 0; j < nclass; j++)
                    if (mj < 0 || yb[mj] < yb[j]) mj = j;
                if (mj == labels.get1(i0 + b)) ncorrect++;
                if (log && ((i0 + b) % 1000) == 0) fprintf(stderr, "i=%d\n", i0 + b);
            }
        }
        return ncorrect;
    }
    for (int i = 0; i < ntests; i++) {
        const uint8_t* img = images.get3(i);
        for (size_t j = 0; j < x.size(); j++) x[j] = img[j] / 255.0;
        net.set_inputs(x.data());
        net.get_outputs(y.data());
        int mj = -1;
        for (int j = 0; j < nclass; j++)
            if (mj < 0 || y[mj] < y[j]) mj = j;
        if (mj == labels.get1(i)) ncorrect++;
        if (log && (i % 1000) == 0) fprintf(stderr, "i=%d\n", i);
    }
    return ncorrect;
}

int run_serial(const TrainOptions& opt) {
    GlibcRand rng(opt.seed);
    Net net = build_net(opt, rng);
    DataSet data;
    if (int rc = load_data(opt, data)) return rc;
    std::string err;
    if (!opt.load_path.empty() && !net.load(opt.load_path, &err)) {
        fprintf(stderr, "%s\n", err.c_str());
        return kFileError;
    }

    fprintf(stderr, "training...\n");
    const int train_size = int(data.train_x.count());
    const int nclass = net.layer(net.num_layers() - 1).nnodes();
    std::vector<double> x(data.train_x.item_size()), y(nclass);
    const long total = long(opt.nepoch) * train_size;
    double etotal = 0;
    int pending = 0, since_log = 0;
    const double t0 = now_s();
    const bool batched = !opt.ref_quirks && !opt.per_sample && net.batch_supported();
    if (batched) {
        // Same samples, same order, same updates as the loop below; the mini-batch goes through Net::learn_batch.
        const int B = opt.batch_size;
        const size_t item = data.train_x.item_size();
        std::vector<double> xs(size_t(B) * item), ys(size_t(B) * nclass), errs(B);
        for (long i = 0; i < total;) {
            const int nb = int(std::min<long>(B, total - i));
            for (int b = 0; b < nb; b++) {
                const int index = rng.next() % train_size;
                load_sample(data.train_x, data.train_y, index, &xs[b * item], &ys[size_t(b) * nclass], nclass);
            }
            net.learn_batch(xs.data(), ys.data(), nb, errs.data());
            for (int b = 0; b < nb; b++) {
                etotal += errs[b];
                since_log++;
                if (((i + b) % opt.log_every) == 0) {
                    fprintf(stderr, "i=%ld, error=%.4f\n", i + b, etotal / since_log);
                    etotal = 0;
                    since_log = 0;
                }
            }
            net.update(opt.rate / nb);
            i += nb;
        }
    }
    for (long i = 0; !batched && i < total; i++) {
        // Sampling with replacement from the same stream that initialised the weights (cnn.c:455).
        const int index = rng.next() % train_size;
        load_sample(data.train_x, data.train_y, index, x.data(), y.data(), nclass);
        net.set_inputs(x.data());
        net.learn_outputs(y.data());
        etotal += net.error_total();
        pending++;
        since_log++;
        if (opt.ref_quirks) {
            if ((i % opt.batch_size) == 0) {  // D2, cnn.c:467
                net.update(opt.rate / opt.batch_size);
                pending = 0;
            }
            if ((i % opt.log_every) == 0) {
                fprintf(stderr, "i=%ld, error=%.4f\n", i, etotal / opt.log_every);
                etotal = 0;
            }
        } else {
            if (pending == opt.batch_size) {
                net.update(opt.rate / opt.batch_size);
                pending = 0;
            }
            if ((i % opt.log_every) == 0) {
                fprintf(stderr, "i=%ld, error=%.4f\n", i, etotal / since_log);
                etotal = 0;
                since_log = 0;
            }
        }
    }
    if (!opt.ref_quirks && pending > 0) net.update(opt.rate / pending);  // tail batch
    const double t1 = now_s();
    if (opt.timing && total > 0)
        fprintf(stderr, "train_images_per_sec=%.1f (%ld samples in %.3f s)\n", total / (t1 - t0), total, t1 - t0);

    if (!opt.save_path.empty() && !net.save(opt.save_path, &err)) {
        fprintf(stderr, "%s\n", err.c_str());
        return kFileError;
    }
    fprintf(stderr, "testing...\n");
    const int ncorrect = evaluate(net, data.test_x, data.test_y, true);
    fprintf(stderr, "ntests=%d, ncorrect=%d\n", int(data.test_x.count()), ncorrect);
    return kOk;
}

int run_data_parallel(const TrainOptions& opt, Comm& comm) {
    const int rank = comm.rank(), ws = comm.size();
    // Every rank seeds identically AND rank 0's parameters are broadcast, so replicas start
    // equal (the reference seeds per rank and never synchronises: SURVEY D7).
    GlibcRand rng(opt.seed);
    Net net = build_net(opt, rng);
    DataSet data;
    int rc = load_data(opt, data);
    // A rank that cannot read its files must not leave the others hanging in a collective:
    // agree on the worst status first.
    double status = rc;
    comm.allreduce_sum(&status, 1);
    if (status != 0) return rc ? rc : kFileError;
    std::string err;
    // Same agreement for the checkpoint: a failed --load on rank 0 ends the run on every rank (run_serial returns
    // kFileError too) instead of silently training from the random initialisation.
    status = 0;
    if (!opt.load_path.empty() && rank == 0 && !net.load(opt.load_path, &err)) {
        fprintf(stderr, "%s\n", err.c_str());
        status = 1;
    }
    comm.allreduce_sum(&status, 1);
    if (status != 0) return kFileError;

    std::vector<double> flat;
    net.gather_params(flat);
    comm.broadcast(flat.data(), flat.size(), 0);
    net.scatter_params(flat);

    const int train_size = int(data.train_x.count());
    const int shard = train_size / ws;  // remainder dropped, cnnmpi.c:457-458
    const int start = shard * rank, end = shard * (rank + 1);
    fprintf(stderr, "%d %d %d\n", rank, start, end);  // cnnmpi.c:459
    if (rank == 0) fprintf(stderr, "training...\n");

    const int nclass = net.layer(net.num_layers() - 1).nnodes();
    std::vector<double> x(data.train_x.item_size()), y(nclass);
    double etotal = 0;
    int since_log = 0;
    comm.barrier();
    const double t0 = now_s();
    for (int epoch = 0; epoch < opt.nepoch; epoch++) {
        if (rank == 0) fprintf(stderr, "epoch = %d\n", epoch);
        int pending = 0;
        if (!opt.per_sample && net.batch_supported()) {
            const int B = opt.batch_size;
            const size_t item = data.train_x.item_size();
            std::vector<double> xs(size_t(B) * item), ys(size_t(B) * nclass), errs(B);
            for (int idx = start; idx < end;) {
                const int nb = std::min(B, end - idx);
                for (int b = 0; b < nb; b++)
                    load_sample(data.train_x, data.train_y, idx + b, &xs[b * item], &ys[size_t(b) * nclass], nclass);
                net.learn_batch(xs.data(), ys.data(), nb, errs.data());
                net.gather_updates(flat);
                comm.allreduce_sum(flat.data(), flat.size());
                net.scatter_updates(flat);
                net.update(opt.rate / (double(nb) * ws));
                for (int b = 0; b < nb; b++) {
                    etotal += errs[b];
                    since_log++;
                    if (rank == 0 && ((idx + b - start) % opt.log_every) == 0) {
                        fprintf(stderr, "    idx = %d, error = %f\n", idx + b, etotal / since_log);
                        etotal = 0;
                        since_log = 0;
                    }
                }
                idx += nb;
            }
            continue;
        }
        for (int idx = start; idx < end; idx++) {
            load_sample(data.train_x, data.train_y, idx, x.data(), y.data(), nclass);
            net.set_inputs(x.data());
            net.learn_outputs(y.data());
            etotal += net.error_total();
            since_log++;
            pending++;
            if (pending == opt.batch_size || idx == end - 1) {
                // The intended algorithm of cnnmpi.c:487-499: sum the weight / bias gradients of
                // the mini-batch over ranks, average, SGD step.
                net.gather_updates(flat);
                comm.allreduce_sum(flat.data(), flat.size());
                net.scatter_updates(flat);
                net.update(opt.rate / (double(pending) * ws));
                pending = 0;
            }
            if (rank == 0 && ((idx - start) % opt.log_every) == 0) {
                fprintf(stderr, "    idx = %d, error = %f\n", idx, etotal / since_log);
                etotal = 0;
                since_log = 0;
            }
        }
    }
    comm.barrier();
    const double t1 = now_s();
    if (rank == 0 && opt.timing) {
        const long total = long(opt.nepoch) * shard * ws;
        fprintf(stderr, "train_images_per_sec=%.1f (%ld samples, %d ranks, %.3f s)\n", total / (t1 - t0), total, ws,
                t1 - t0);
    }
    if (rank == 0) {
        if (!opt.save_path.empty() && !net.save(opt.save_path, &err)) fprintf(stderr, "%s\n", err.c_str());
        fprintf(stderr, "testing...\n");
        const int ncorrect = evaluate(net, data.test_x, data.test_y, true);
        fprintf(stderr, "ntests=%d, ncorrect=%d\n", int(data.test_x.count()), ncorrect);
    }
    // Replica-consistency check value: every rank prints a checksum of its parameters.
    net.gather_params(flat);
    double sum = 0;
    for (double v : flat) sum += v;
    fprintf(stderr, "rank %d param_checksum=%.17g\n", rank, sum);
    comm.barrier();
    return kOk;
}

}  // namespace b200cnn
