// Error channel + version string of libvllm_b200.
#include <stdarg.h>

#include "common.cuh"

namespace vb200 {
static thread_local char g_err[1024] = {0};
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace vb200

extern "C" const char* vb200_last_error(void) { return vb200::g_err; }
extern "C" const char* vb200_version(void) { return "vllm_b200 0.1 sm_100a"; }
