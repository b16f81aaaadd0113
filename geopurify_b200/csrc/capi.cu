// Error state and version of libgeopurify_b200.so.
#include <stdarg.h>
#include "common.cuh"

static thread_local char g_err[512] = "";

void gp_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* gp_last_error(void) { return g_err; }
extern "C" int gp_version(void) { return 100; }
