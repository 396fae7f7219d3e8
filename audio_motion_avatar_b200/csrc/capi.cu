// Error string + version of the C ABI (include/amav.h).
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void amav_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *amav_last_error(void) { return g_err; }
extern "C" int amav_version(void) { return 100; }

// Kernel launches issued by this library since load (bench.py reports the per-step delta as gpu_launches).
static unsigned long long g_launches = 0;
void amav_count_launches(int n) { __atomic_fetch_add(&g_launches, (unsigned long long)n, __ATOMIC_RELAXED); }
extern "C" int64_t amav_launch_count(void) { return (int64_t)__atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-(function, device) attribute: set it when a launch needs more
// than what was set before on that device, not on every call (the entry points sit on per-frame-batch hot paths and
// inside CUDA-graph captures).
#include <mutex>
#include <vector>
namespace {
struct SmemOptIn { const void *func; int device; int bytes; };
std::mutex g_smem_mu;
std::vector<SmemOptIn> g_smem;
}
cudaError_t amav_ensure_dyn_smem(const void *func, int bytes)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(g_smem_mu);
    for (auto &r : g_smem)
        if (r.func == func && r.device == dev) {
            if (r.bytes >= bytes) return cudaSuccess;
            e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
            if (e == cudaSuccess) r.bytes = bytes;
            return e;
        }
    e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) g_smem.push_back({func, dev, bytes});
    return e;
}
