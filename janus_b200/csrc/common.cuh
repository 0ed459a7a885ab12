// Shared host/device helpers for libjanus_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/janus_b200.h"

namespace janus {

// thread-local last-error buffer (defined in abi.cu)
void set_error(const char* fmt, ...);

#define JANUS_CHECK_ARG(cond, ...)                 \
  do {                                             \
    if (!(cond)) {                                 \
      ::janus::set_error(__VA_ARGS__);             \
      return JANUS_ERR_INVALID_ARG;                \
    }                                              \
  } while (0)

#define JANUS_CUDA(call)                                                              \
  do {                                                                                \
    cudaError_t e__ = (call);                                                         \
    if (e__ != cudaSuccess) {                                                         \
      ::janus::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),    \
                         __FILE__, __LINE__);                                         \
      return JANUS_ERR_CUDA;                                                          \
    }                                                                                 \
  } while (0)

#define JANUS_LAUNCH_CHECK(name)                                                      \
  do {                                                                                \
    cudaError_t e__ = cudaGetLastError();                                             \
    if (e__ != cudaSuccess) {                                                         \
      ::janus::set_error("launch of %s failed: %s", name, cudaGetErrorString(e__));  \
      return JANUS_ERR_CUDA;                                                          \
    }                                                                                 \
  } while (0)

constexpr unsigned FULL_MASK = 0xffffffffu;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL_MASK, v, o));
  return v;
}

}  // namespace janus
