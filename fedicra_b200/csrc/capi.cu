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
