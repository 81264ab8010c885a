// mkidx: writes a synthetic MNIST-shaped data set as four IDX files (replaces the network
// download of /root/reference/Makefile:24-35, which cannot work offline).
//   mkidx OUTDIR [ntrain=1000] [ntest=200]
#include <cstdio>
#include <cstdlib>
#include <string>

#include "idx.hpp"

int main(int argc, char** argv) {
    if (argc < 2) {
        fprintf(stderr, "usage: %s OUTDIR [ntrain] [ntest]\n", argv[0]);
        return 100;
    }
    const std::string dir = argv[1];
    const unsigned ntrain = argc > 2 ? unsigned(atoi(argv[2])) : 1000;
    const unsigned ntest = argc > 3 ? unsigned(atoi(argv[3])) : 200;
    b200cnn::IdxFile x, y;
    std::string err;
    b200cnn::make_synthetic(ntrain, 28, 28, 10, 1, x, y);
    if (!b200cnn::idx_write(dir + "/train-images-idx3-ubyte", x, &err) ||
        !b200cnn::idx_write(dir + "/train-labels-idx1-ubyte", y, &err)) {
        fprintf(stderr, "%s\n", err.c_str());
        return 111;
    }
    b200cnn::make_synthetic(ntest, 28, 28, 10, 2, x, y);
    if (!b200cnn::idx_write(dir + "/t10k-images-idx3-ubyte", x, &err) ||
        !b200cnn::idx_write(dir + "/t10k-labels-idx1-ubyte", y, &err)) {
        fprintf(stderr, "%s\n", err.c_str());
        return 111;
    }
    return 0;
}
