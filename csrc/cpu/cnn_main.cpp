// Serial CPU trainer entrypoint (configuration 1; replaces /root/reference/cnn.c:406-531).
#include "trainer.hpp"

int main(int argc, char** argv) {
    b200cnn::TrainOptions opt;
    if (int rc = b200cnn::parse_options(argc, argv, /*dp=*/false, opt)) return rc;
    return b200cnn::run_serial(opt);
}
