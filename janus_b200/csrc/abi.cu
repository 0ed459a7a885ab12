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
// process-wide tunables (janus_set_option): -1 = "use the built-in default" at the site that reads the option
static int g_options[JANUS_OPT_COUNT] = {-1, -1, -1, -1, -1, -1};
int option(int opt, int dflt) {
  const int v = (opt >= 0 && opt < JANUS_OPT_COUNT) ? g_options[opt] : -1;
  return v < 0 ? dflt : v;
}
}  // namespace janus

extern "C" int32_t janus_set_option(int32_t opt, int32_t value) {
  if (opt < 0 || opt >= JANUS_OPT_COUNT) {
    janus::set_error("janus_set_option: unknown option %d", opt);
    return JANUS_ERR_INVALID_ARG;
  }
  janus::g_options[opt] = value < 0 ? -1 : value;
  if (opt == JANUS_OPT_PDL) janus::g_pdl_enabled = value != 0;   // default (-1) = on
  return JANUS_OK;
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
