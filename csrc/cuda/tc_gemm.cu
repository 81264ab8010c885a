// placeholder until the tcgen05 path lands: nothing is supported, the SIMT path runs.
#include "common.cuh"
#include "backend.cuh"
namespace b200cnn {
bool tc_linear_fwd_supported(int, int, int, int) { return false; }
bool tc_linear_bwd_data_supported(int, int, int, int) { return false; }
bool tc_linear_bwd_weight_supported(int, int, int, int) { return false; }
int tc_linear_fwd(const float*, const float*, const float*, float*, int, int, int, int, int, cudaStream_t) { return set_error(kInvalidValue, "tcgen05 path not built"); }
int tc_linear_bwd_data(const float*, const float*, const float*, int, float*, int, int, int, int, cudaStream_t) { return set_error(kInvalidValue, "tcgen05 path not built"); }
int tc_linear_bwd_weight(const float*, const float*, float*, float*, int, int, int, int, int, cudaStream_t) { return set_error(kInvalidValue, "tcgen05 path not built"); }
}
