// Thread-local last-error string behind the C ABI (see include/tfdl_kernels.h).
#include <cstdarg>
#include <cstdio>
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {
static thread_local char g_err[1024] = "";
void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace tfdl

extern "C" const char* tfdl_last_error(void) { return tfdl::g_err; }
extern "C" int tfdl_kernels_abi_version(void) { return 1; }
