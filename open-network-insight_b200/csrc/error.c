/* Per-thread error message behind onilda_last_error(). */
#include <stdarg.h>
#include <stdio.h>

#include "onilda_internal.h"

static __thread char g_err[512] = "";

int onilda_fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

const char* onilda_last_error(void) { return g_err; }
const char* onilda_version(void) { return "onilda-b200 0.1 (sm_100a)"; }
