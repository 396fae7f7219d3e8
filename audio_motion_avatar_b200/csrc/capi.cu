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
