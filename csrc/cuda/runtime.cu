// Error reporting and build facts of the GPU library.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "common.cuh"

namespace b200cnn {

static thread_local char g_error[512] = "";

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
    return code ? code : kInvalidValue;
}

bool pdl_enabled() {
    static const bool on = [] {
        const char* v = getenv("B200CNN_PDL");
        return !(v && v[0] == '0');
    }();
    return on;
}

static thread_local cudaStream_t g_critical_stream = nullptr;
cudaStream_t critical_stream() { return g_critical_stream; }
int critical_priority() {
    static const int prio = [] {
        const char* v = getenv("B200CNN_PRIORITY");
        if (v && v[0] == '0') return 0;
        int least = 0, greatest = 0;
        if (cudaDeviceGetStreamPriorityRange(&least, &greatest) != cudaSuccess) return 0;
        return greatest;  // numerically lowest = most urgent; 0 when the device has a single level
    }();
    return prio;
}
CriticalStreamScope::CriticalStreamScope(cudaStream_t st) : saved(g_critical_stream) { g_critical_stream = st; }
CriticalStreamScope::~CriticalStreamScope() { g_critical_stream = saved; }

}  // namespace b200cnn

extern "C" const char* b200cnn_last_error(void) { return b200cnn::g_error; }

extern "C" const char* b200cnn_build_info(void) {
#define B200CNN_STR2(x) #x
#define B200CNN_STR(x) B200CNN_STR2(x)
    return "b200cnn cuda library; target sm_100a; CUDA " B200CNN_STR(CUDART_VERSION) "; features: "
#ifdef B200CNN_HAVE_TCGEN05
           "tcgen05-gemm tma "
#endif
           "simt p2p-allreduce-sgd nvls";
}
