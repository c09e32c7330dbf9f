// Error channel + version string of libvllm_b200.
#include <stdarg.h>
#include <stdlib.h>

#include "common.cuh"

namespace vb200 {
static thread_local char g_err[1024] = {0};
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("VB200_PDL");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}
}  // namespace vb200

extern "C" const char* vb200_last_error(void) { return vb200::g_err; }
extern "C" const char* vb200_version(void) { return "vllm_b200 0.2 sm_100a"; }
