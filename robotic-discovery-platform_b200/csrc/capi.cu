// Error reporting and version entry points of the C ABI (include/rdp_b200.h).
#include <stdarg.h>
#include <stdio.h>

#include "rdp_common.cuh"

namespace rdp {
static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}
}  // namespace rdp

extern "C" int rdp_abi_version(void) { return 3; }
extern "C" const char* rdp_last_error(void) { return rdp::g_error; }
