// C-ABI status plumbing: ABI version and the per-thread last-error message.
#include <stdarg.h>
#include <stdio.h>

#include "common.cuh"
#include "puflow_b200.h"

static thread_local char g_err[512] = "";

void pf_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" int pf_abi_version(void) { return PF_ABI_VERSION; }
extern "C" const char* pf_last_error(void) { return g_err; }
