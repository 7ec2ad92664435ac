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

// Launch shape of the memory-bound helper kernels (pack + norm, sum of squares, AdamW).  "Small footprint" = 128-thread
// blocks, at most two per SM: such a block (<= 6 K registers, no shared memory to speak of) fits beside a resident
// persistent GEMM / attention CTA, so helper work issued on a side stream really runs UNDER the backward pass instead of
// taking turns with it at kernel boundaries.  Process-wide; read at launch (and therefore baked into a captured graph).
static int g_small_footprint = 0;
extern "C" void tdl_set_small_footprint(int on) { g_small_footprint = on ? 1 : 0; }
int tdl_small_footprint() { return g_small_footprint; }
