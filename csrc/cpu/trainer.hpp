// Shared training / evaluation driver for the two CPU entrypoints.
//
//   cnn      = the reference's serial trainer    (/root/reference/cnn.c:406-531)
//   cnnmpi   = the reference's MPI CPU trainer   (/root/reference/cnnmpi.c:412-560)
//
// Command line kept from the reference: four positional IDX paths
//   prog train-images train-labels test-images test-labels          (cnn.c:408-411)
// The reference's hyper-parameters are literals (cnn.c:446-449, :413); here they are optional
// flags whose defaults are those literals.
#pragma once
#include <string>

#include "comm.hpp"
#include "idx.hpp"
#include "net.hpp"

namespace b200cnn {

struct TrainOptions {
    std::string train_images, train_labels, test_images, test_labels;
    double rate = 0.1;       // cnn.c:446
    int nepoch = 10;         // cnn.c:448
    int batch_size = 32;     // cnn.c:449
    unsigned seed = 0;       // cnn.c:413
    int log_every = 1000;    // cnn.c:470
    int synthetic = 0;       // >0: generate this many training samples instead of reading files
    int synthetic_test = 0;  // test samples for --synthetic (default: synthetic/5, at least 100)
    bool ref_quirks = false;  // reproduce D1/D2/D4 for diffing against the unmodified cnn.c
    bool per_sample = false;  // clean mode: keep the one-sample-at-a-time schedule (A/B against Net::learn_batch)
    bool pool = false;       // conv -> relu -> 2x2 max-pool variant of the topology
    std::string save_path, load_path;
    int np = 0;              // cnnmpi only: ranks to fork (0 = take RANK/WORLD_SIZE from the env)
    bool timing = false;     // print images/s on stderr after training
};

// Exit codes of the reference: 100 = too few arguments (cnn.c:412), 111 = file error
// (cnn.c:432-442), 0 = success.
enum ExitCode { kOk = 0, kUsage = 100, kFileError = 111 };

// Parses argv.  Returns kOk or kUsage (after printing a usage line on stderr).
int parse_options(int argc, char** argv, bool dp, TrainOptions& opt);

struct DataSet {
    IdxFile train_x, train_y, test_x, test_y;
};
int load_data(const TrainOptions& opt, DataSet& data);

Net build_net(const TrainOptions& opt, GlibcRand& rng);

// Serial training loop + test pass; prints the reference's stderr lines.
int run_serial(const TrainOptions& opt);
// Data-parallel training over `comm`; rank 0 evaluates (cnnmpi.c:521).
int run_data_parallel(const TrainOptions& opt, Comm& comm);

// Forward every test image, first-max arg-max (cnn.c:508-513); returns #correct.
int evaluate(Net& net, const IdxFile& images, const IdxFile& labels, bool log);

}  // namespace b200cnn
