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
// process-wide tunable `opt` (janus_option_t), or `dflt` when it was never set / reset to -1 (abi.cu)
int option(int opt, int dflt);
// index of the current device for per-device caches of kernel attributes (gemm.cu)
int current_device_slot();

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

// ---- programmatic dependent launch (PDL) ------------------------------------------------------
// The decode step is a chain of ~140 tiny dependent kernels; each one is launched with
// programmaticStreamSerialization so that its CTAs become resident and prefetch their immutable
// operands (weights, LayerNorm parameters, cross-attention keys) while the producer kernel is still
// running. Contract for every kernel in such a chain:
//   * before pdl_wait(): only reads of data that no kernel of the chain writes; no writes at all
//   * after  pdl_wait(): the producer grid (and, transitively, every earlier grid) has completed and
//     its writes are visible; mutable data is read with ordinary loads (never ld.global.nc)
// Launched without the attribute both instructions are no-ops.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

extern int g_pdl_enabled;  // janus_set_pdl(): 0 launches the same kernels fully serialised (A/B measurements)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = g_pdl_enabled ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

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
