// Error text, launch counter and version for the C ABI (include/cnmf_b200.h).
#include <stdarg.h>

#include "common.cuh"

namespace cnmf {

static thread_local char g_error[512] = "";
static thread_local int64_t g_launches = 0;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}
void count_launch(int64_t n) { g_launches += n; }
void reset_launch_count() { g_launches = 0; }

}  // namespace cnmf

extern "C" {
int cnmf_b200_abi_version(void) { return CNMF_B200_ABI_VERSION; }
const char* cnmf_b200_last_error(void) { return cnmf::g_error; }
int64_t cnmf_b200_last_launch_count(void) { return cnmf::g_launches; }
}
