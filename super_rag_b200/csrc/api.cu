// Library-level entry points: ABI version, thread-local error text, device info.
#include <stdarg.h>
#include <string.h>

#include <map>
#include <mutex>
#include <utility>

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
    return srag_current_device_props(sm_count, cc_major, cc_minor);
}


// ---------------------------------------------------------------------------------------
// Per-device state. cudaFuncSetAttribute(MaxDynamicSharedMemorySize) and the SM count belong to a DEVICE, and one
// process may drive several (ctx["device"] = "cuda:1"): both caches are keyed by the current device ordinal and
// guarded by a mutex (the C ABI may be entered from several host threads).
// ---------------------------------------------------------------------------------------
namespace {
std::mutex g_dev_mu;
std::map<std::pair<const void*, int>, int> g_smem_attr;  // (kernel, device) -> bytes opted in
struct DevProps { int sms, major, minor; };
std::map<int, DevProps> g_dev_props;
}  // namespace

int srag_func_smem_optin(const void* func, int bytes) {
    int dev = 0;
    SRAG_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_dev_mu);
    auto it = g_smem_attr.find({func, dev});
    if (it != g_smem_attr.end() && it->second >= bytes) return SRAG_OK;
    SRAG_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    g_smem_attr[{func, dev}] = bytes;
    return SRAG_OK;
}

int srag_current_device_props(int* sms, int* major, int* minor) {
    int dev = 0;
    SRAG_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_dev_mu);
    auto it = g_dev_props.find(dev);
    if (it == g_dev_props.end()) {
        DevProps p;
        SRAG_CUDA(cudaDeviceGetAttribute(&p.sms, cudaDevAttrMultiProcessorCount, dev));
        SRAG_CUDA(cudaDeviceGetAttribute(&p.major, cudaDevAttrComputeCapabilityMajor, dev));
        SRAG_CUDA(cudaDeviceGetAttribute(&p.minor, cudaDevAttrComputeCapabilityMinor, dev));
        it = g_dev_props.emplace(dev, p).first;
    }
    if (sms) *sms = it->second.sms;
    if (major) *major = it->second.major;
    if (minor) *minor = it->second.minor;
    return SRAG_OK;
}
