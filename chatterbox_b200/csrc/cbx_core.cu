// Error plumbing and version of the chatterbox_b200 C-ABI (include/cbx.h).
#include "common.cuh"
#include "../../include/cbx.h"
#include <stdarg.h>

static thread_local char g_err[1024] = "";

void cbx_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* cbx_last_error(void) { return g_err; }
extern "C" int cbx_version(void) { return 1; }
