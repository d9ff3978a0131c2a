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
static unsigned long long g_launches = 0;
void note_launches(int n) { __atomic_fetch_add(&g_launches, (unsigned long long)n, __ATOMIC_RELAXED); }
}  // namespace dpvo

extern "C" int dpvo_abi_version(void) { return DPVO_B200_ABI_VERSION; }
extern "C" const char* dpvo_last_error(void) { return dpvo::g_err; }
extern "C" uint64_t dpvo_launch_count(void) { return __atomic_load_n(&dpvo::g_launches, __ATOMIC_RELAXED); }
