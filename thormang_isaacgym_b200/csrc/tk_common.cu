// Error plumbing and device queries shared by the task translation units.
#include "tk_common.cuh"

#include <cstdarg>
#include <cstdio>

namespace tk {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) in %s", static_cast<int>(e), cudaGetErrorString(e), what);
    return TK_ERR_CUDA;
}

const char* last_error() { return g_error; }

__global__ void advance_epoch_kernel(unsigned long long* epoch) { *epoch += 1ull; }

}  // namespace tk

extern "C" {

const char* tk_last_error(void) { return tk::last_error(); }

int tk_abi_version(void) { return TK_ABI_VERSION; }

int tk_device_info(int device, int* sm_count, int* cc_major, int* cc_minor) {
    cudaDeviceProp prop;
    TK_CUDA(cudaGetDeviceProperties(&prop, device));
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    return TK_OK;
}

}  // extern "C"
