// Internal helpers shared by the translation units of libs2p_sm100.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include "../../include/s2p_b200.h"

namespace s2p {

// last error text, per host thread (returned by s2p_last_error_string()).
char* error_buffer();
int set_error(int code, const char* fmt, ...);

#define S2P_CHECK_ARG(cond, ...)                                   \
  do {                                                             \
    if (!(cond)) return ::s2p::set_error(S2P_ERR_BAD_ARG, __VA_ARGS__); \
  } while (0)

#define S2P_CHECK_CUDA(expr)                                                          \
  do {                                                                                \
    cudaError_t e__ = (expr);                                                         \
    if (e__ != cudaSuccess)                                                           \
      return ::s2p::set_error(S2P_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e__)); \
  } while (0)

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(S2P_ERR_CUDA, "%s launch failed: %s", what, cudaGetErrorString(e));
  return S2P_OK;
}

// Per-device caches (SM count, cudaFuncSetAttribute done-flags) are indexed by the CURRENT device's ordinal:
// a process that drives a second GPU must not inherit the first one's state.
constexpr int kMaxDevices = 64;
int current_device();   // ordinal of the calling thread's current device, clamped to [0, kMaxDevices)
int num_sms();          // SM count of the current device
void count_launch();

}  // namespace s2p
