// capi.cu -- error text and version of the C ABI.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

int fedicra_fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

extern "C" const char *fedicra_last_error(void) { return g_err; }
extern "C" int fedicra_abi_version(void) { return FEDICRA_ABI_VERSION; }

static long long g_launches = 0;
void fedicra_count_launches(int n) { __atomic_fetch_add(&g_launches, (long long)n, __ATOMIC_RELAXED); }
extern "C" long long fedicra_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }
