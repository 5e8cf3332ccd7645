// Host-side error plumbing shared by every C-ABI entry point.
#include "common.cuh"

#include <cstdarg>
#include <cstdio>

namespace sq {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return SQ_ERR_CUDA;
}

int device_sm_count() {
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 148;
    return n;
}

}  // namespace sq

extern "C" const char* sq_last_error(void) { return sq::g_err; }
extern "C" int sq_abi_version(void) { return 1; }
