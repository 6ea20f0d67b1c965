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

int num_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
      sms = 148;
  }
  return sms;
}

}  // namespace s2p

extern "C" int s2p_version(void) { return 1; }
extern "C" const char* s2p_last_error_string(void) { return s2p::error_buffer(); }
extern "C" long long s2p_launch_count(void) { return s2p::g_launches.load(std::memory_order_relaxed); }
