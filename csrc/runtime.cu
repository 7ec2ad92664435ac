// Library-wide runtime bits: error string, version, device query.
#include "common.cuh"
#include <cstdarg>
#include <cstdio>

static thread_local char g_err[1024] = "";

int tdl_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}

extern "C" const char* tdl_last_error(void) { return g_err; }
extern "C" int tdl_version(void) { return 100; }
extern "C" int tdl_device_sm_count(void) { return tdl_num_sms(); }

