// Error plumbing + ABI version for libdpvo_b200.
#include "common.cuh"
#include <stdarg.h>

namespace dpvo {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace dpvo

extern "C" int dpvo_abi_version(void) { return DPVO_B200_ABI_VERSION; }
extern "C" const char* dpvo_last_error(void) { return dpvo::g_err; }
