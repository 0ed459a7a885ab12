// LayerNorm over the last dimension, bf16 in -> bf16 out, fp32 statistics (eps 1e-5), one warp per
// row. Memory-bound (reads + writes d*2 B per row once); used between the tensor-core GEMMs of
// the Whisper encoder/decoder blocks (pre-LN architecture).
#include <cuda_bf16.h>

#include "kernels.cuh"

namespace janus {

constexpr int LN_MAX_PAIRS = 20;  // d <= 1280 (large-v3)

// x / y carry no __restrict__: under PDL (see common.cuh) x is written by the producer grid while this
// grid is already resident, so it must not be read through the non-coherent path.
__global__ void __launch_bounds__(256) layernorm_bf16_kernel(const __nv_bfloat16* x,
                                                             const float* __restrict__ w,
                                                             const float* __restrict__ b,
                                                             __nv_bfloat16* y, long long rows, int d,
                                                             long long ldx, long long ldy) {
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  pdl_launch_dependents();
  pdl_wait();
  if (row >= rows) return;
  const __nv_bfloat162* xr = reinterpret_cast<const __nv_bfloat162*>(x + row * ldx);
  const int npairs = d >> 6;  // bf16x2 pairs per lane
  float2 v[LN_MAX_PAIRS];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAX_PAIRS; ++i) {
    if (i < npairs) {
      v[i] = __bfloat1622float2(xr[lane + 32 * i]);
      sum += v[i].x + v[i].y;
    }
  }
  const float mean = warp_sum(sum) / (float)d;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAX_PAIRS; ++i) {
    if (i < npairs) {
      const float a = v[i].x - mean, c = v[i].y - mean;
      sq += a * a + c * c;
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / (float)d + 1e-5f);
  __nv_bfloat162* yr = reinterpret_cast<__nv_bfloat162*>(y + row * ldy);
  const float2* w2 = reinterpret_cast<const float2*>(w);
  const float2* b2 = reinterpret_cast<const float2*>(b);
#pragma unroll
  for (int i = 0; i < LN_MAX_PAIRS; ++i) {
    if (i < npairs) {
      const float2 ww = __ldg(w2 + lane + 32 * i), bb = __ldg(b2 + lane + 32 * i);
      yr[lane + 32 * i] = __floats2bfloat162_rn((v[i].x - mean) * rstd * ww.x + bb.x,
                                                (v[i].y - mean) * rstd * ww.y + bb.y);
    }
  }
}

int layernorm_bf16(const __nv_bfloat16* x, const float* w, const float* b, __nv_bfloat16* y, long long rows, int d,
                   long long ldx, long long ldy, cudaStream_t stream, bool pdl) {
  JANUS_CHECK_ARG(d % 64 == 0 && d <= 64 * LN_MAX_PAIRS, "layernorm: d=%d must be a multiple of 64 and <= %d", d,
                  64 * LN_MAX_PAIRS);
  if (rows <= 0) return JANUS_OK;
  const int wpb = 8;
  const unsigned grid = (unsigned)((rows + wpb - 1) / wpb);
  if (pdl) {
    JANUS_CUDA(launch_pdl(layernorm_bf16_kernel, dim3(grid), dim3(wpb * 32), 0, stream, x, w, b, y, rows, d, ldx, ldy));
    return JANUS_OK;
  }
  layernorm_bf16_kernel<<<grid, wpb * 32, 0, stream>>>(x, w, b, y, rows, d, ldx, ldy);
  JANUS_LAUNCH_CHECK("layernorm_bf16_kernel");
  return JANUS_OK;
}

}  // namespace janus

extern "C" int32_t janus_layernorm_bf16(const void* x, const float* w, const float* b, void* y, int64_t rows,
                                        int32_t d, void* stream) {
  JANUS_CHECK_ARG(x && w && b && y, "layernorm: null pointer");
  return janus::layernorm_bf16(static_cast<const __nv_bfloat16*>(x), w, b, static_cast<__nv_bfloat16*>(y), rows, d, d,
                               d, (cudaStream_t)stream);
}
