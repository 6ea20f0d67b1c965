// Library-wide plumbing of libs2p_sm100.so: error text, launch counter, device info.
#include <atomic>

#include "s2p_internal.h"

namespace s2p {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

char* error_buffer() { return g_err; }

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) dev = 0;
  return dev < kMaxDevices ? dev : kMaxDevices - 1;
}

int num_sms() {
  static int sms[kMaxDevices] = {};
  const int dev = current_device();
  if (sms[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    sms[dev] = v;
  }
  return sms[dev];
}

}  // namespace s2p

extern "C" int s2p_version(void) { return 1; }
extern "C" const char* s2p_last_error_string(void) { return s2p::error_buffer(); }
extern "C" long long s2p_launch_count(void) { return s2p::g_launches.load(std::memory_order_relaxed); }
