// Library-level entry points: ABI version, thread-local error text, device info.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void srag_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" int srag_abi_version(void) { return SRAG_ABI_VERSION; }

extern "C" const char* srag_last_error(void) { return g_err; }

extern "C" int srag_device_info(int* sm_count, int* cc_major, int* cc_minor) {
    int dev = 0;
    SRAG_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    SRAG_CUDA(cudaGetDeviceProperties(&prop, dev));
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    return SRAG_OK;
}
