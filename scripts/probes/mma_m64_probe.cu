// Probe: where do the 64 rows of an M=64, cta_group::1 tcgen05.mma accumulator live in TMEM?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I janus_b200/csrc scripts/probes/mma_m64_probe.cu \
//        janus_b200/csrc/abi.cu -o gpurun_out/m64probe   (abi.cu provides g_pdl_enabled / error state)
#include "tc_common.cuh"
#include <cstdio>
using namespace janus;

__global__ void probe(float* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;            // [64 rows][64 cols] bf16, 128B-swizzled, K-major
  uint8_t* sB = smem + 8192;     // [16 rows][64 cols]
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 8192 + 2048);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (8192 + 2048) / 2; i += blockDim.x) reinterpret_cast<__nv_bfloat16*>(smem)[i] = __float2bfloat16(0.f);
  __syncthreads();
  if (threadIdx.x < 64) {   // A[r][0] = r + 1
    const int r = threadIdx.x, c = 0;
    *reinterpret_cast<__nv_bfloat16*>(sA + r * 128 + (((c >> 3) ^ (r & 7)) << 4) + (c & 7) * 2) = __float2bfloat16((float)(r + 1));
  }
  if (threadIdx.x < 16) {   // B[n][0] = n == 0 ? 1 : (n == 1 ? 2 : 0)  -> D[r][0] = r+1, D[r][1] = 2(r+1)
    const int n = threadIdx.x, c = 0;
    *reinterpret_cast<__nv_bfloat16*>(sB + n * 128 + (((c >> 3) ^ (n & 7)) << 4) + (c & 7) * 2) =
        __float2bfloat16(n == 0 ? 1.f : (n == 1 ? 2.f : 0.f));
  }
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  fence_proxy_async_smem();
  if (warp == 0) tmem_alloc(slot, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = *slot;
  // clear the accumulator region first (128 lanes x 16 cols) with -1 so untouched lanes are visible
  {
    uint32_t r[32];
    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(-1.f);
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n" ::
                 "r"(tb + ((uint32_t)(warp * 32) << 16)), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]),
                 "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(64, 16, 0, 0);
      umma_f16(tb, umma_smem_desc_sw128(smem_u32(sA)), umma_smem_desc_sw128(smem_u32(sB)), idesc, 0);
      umma_commit(bar);
    }
    __syncwarp();
  }
  mbar_wait(bar, 0);
  tc_fence_after();
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                 "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(tb + ((uint32_t)(warp * 32) << 16)) : "memory");
  tmem_ld_wait();
  out[(warp * 32 + lane) * 2] = __uint_as_float(r[0]);
  out[(warp * 32 + lane) * 2 + 1] = __uint_as_float(r[1]);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tb, 32); }
}

int main() {
  float* d; cudaMalloc(&d, 128 * 2 * 4);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
  probe<<<1, 128, 16384>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  printf("status %s\n", cudaGetErrorString(e));
  float h[256]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  for (int l = 0; l < 128; ++l) printf("lane %3d: col0 %6.1f col1 %6.1f%s", l, h[2 * l], h[2 * l + 1], (l % 4 == 3) ? "\n" : "   ");
  return 0;
}
