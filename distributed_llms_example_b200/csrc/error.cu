// Thread-local error message behind b2_last_error() (include/b2ddp.h).
#include "common.cuh"

namespace b2 {

static thread_local char g_err[512] = {0};

int fail(int status, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return status;
}

void clear_error() { g_err[0] = 0; }

}  // namespace b2

extern "C" const char* b2_last_error(void) { return b2::g_err; }
extern "C" int b2_abi_version(void) { return B2DDP_ABI_VERSION; }
