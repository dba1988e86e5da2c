// runtime.cu — error reporting and ABI identification for libaffb200.
#include "common.cuh"
#include <stdarg.h>
#include <atomic>

namespace affb200 {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void count_launches(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char *what, int n_kernels) {
    g_launches.fetch_add(n_kernels, std::memory_order_relaxed);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: CUDA error %d (%s)", what, int(e), cudaGetErrorString(e));
        return AFFB200_E_CUDA;
    }
    return AFFB200_OK;
}

}  // namespace affb200

extern "C" int affb200_abi_version(void) { return AFFB200_ABI_VERSION; }
extern "C" const char *affb200_last_error(void) { return affb200::g_err; }
extern "C" const char *affb200_target_arch(void) { return "sm_100a"; }
extern "C" int64_t affb200_launch_count(void) { return (int64_t)affb200::g_launches.load(); }
