// ABI plumbing: version, thread-local error string, device check.
#include "common.cuh"

namespace janus {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int g_pdl_enabled = 1;
}  // namespace janus

extern "C" int32_t janus_set_pdl(int32_t enabled) {
  const int prev = janus::g_pdl_enabled;
  janus::g_pdl_enabled = enabled != 0;
  return prev;
}

extern "C" int32_t janus_abi_version(void) { return JANUS_ABI_VERSION; }
extern "C" const char* janus_last_error(void) { return janus::g_err; }

extern "C" int32_t janus_check_device(int32_t dev) {
  cudaDeviceProp p;
  JANUS_CUDA(cudaGetDeviceProperties(&p, dev));
  if (p.major != 10) {
    janus::set_error("device %d is sm_%d%d; libjanus_b200 contains sm_100a code only", dev, p.major,
                     p.minor);
    return JANUS_ERR_UNSUPPORTED;
  }
  return JANUS_OK;
}
