// Error plumbing and version of the graphnet_b200 C ABI.
#include <stdarg.h>

#include "common.cuh"

namespace gn {
static thread_local char g_error[512] = "";
unsigned long long g_launch_count = 0;
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}
}  // namespace gn

extern "C" const char* gn_last_error(void) { return gn::g_error; }
extern "C" int gn_version(void) { return 100; }
extern "C" uint64_t gn_launch_count(void) { return gn::g_launch_count; }
