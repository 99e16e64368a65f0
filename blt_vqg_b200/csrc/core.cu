// Library-wide state: thread-local error string, ABI version, device check.
#include "common.cuh"
#include <stdarg.h>

namespace b200iq {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static int g_round = 1;
int round_mode() { return g_round; }
static int g_pdl = 0;
int pdl_mode() { return g_pdl; }

int sm_count() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}
}  // namespace b200iq

extern "C" int b200iq_abi_version(void) { return B200IQ_ABI_VERSION; }

extern "C" int b200iq_set_tf32_rounding(int on) {
  int prev = b200iq::g_round;
  b200iq::g_round = on ? 1 : 0;
  return prev;
}

extern "C" int b200iq_set_pdl(int on) {
  int prev = b200iq::g_pdl;
  b200iq::g_pdl = on ? 1 : 0;
  return prev;
}

extern "C" const char* b200iq_last_error(void) { return b200iq::g_err; }

extern "C" int b200iq_device_supported(void) {
  int dev = 0, major = 0, minor = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return 0; }
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  return (major == 10 && minor == 0) ? 1 : 0;
}
