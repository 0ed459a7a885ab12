// Skinny linear layer for the decode step:  out[b, n] = act( sum_k A[b, k] W[n, k] + bias[n] ) (+ residual)
// with b < 64..256 streams. One decode step runs ~73 of these back to back and each touches only
// 1-5 MB of weights exactly once, so the op is bound by launch + one DRAM round trip, not by tensor
// throughput: the tcgen05 GEMM (TMEM allocation, TMA descriptor fetch, mbarrier pipeline fill, 10-warp
// role hand-off) measured 12 us per call here, most of it fixed cost. This kernel has the shortest
// possible dependency chain instead:
//   * every warp streams its 8 weight rows straight from HBM into registers with 16-byte loads
//     (K is permuted identically for A and W, so a thread's 8 consecutive k feed two
//     mma.sync.m16n8k16 directly -- no shared-memory staging of W at all);
//   * the tiny activation matrix is read through L1 (every warp on the SM reads the same lines);
//   * two warps split K for each 8-column tile and combine through shared memory; the epilogue
//     (bias / GELU / residual / bf16 or fp32 store) runs from the accumulator registers.
// mma.sync is the legacy tensor path, used here on purpose: at M <= 256 rows and one pass over the
// weights the 5th-gen tensor core's setup latency is the cost, not its throughput.
#include "kernels.cuh"

namespace janus {

__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

constexpr int DL_COLS = 16;       // output columns per CTA (2 column tiles of 8)

// MT = 16-row tiles of A, KS = number of warps splitting K for each column tile (2 for K <= 1024,
// 8 for the K = 4d MLP down-projection so every warp still streams ~12 32-wide k blocks)
template <int MT, int KS>
__global__ void __launch_bounds__(2 * KS * 32)
dec_linear_kernel(const __nv_bfloat16* A, int lda, const __nv_bfloat16* __restrict__ W,
                  const float* __restrict__ bias, const __nv_bfloat16* residual, void* out,
                  int rows, int N, int K, int act, int out_f32) {
  __shared__ float red[KS - 1][2][MT][32][4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ct = warp & 1, kh = warp >> 1;
  const int g = lane >> 2, t4 = lane & 3;
  const int n0 = blockIdx.x * DL_COLS + ct * 8;
  const int kblocks = K >> 5;
  const int kper = (kblocks + KS - 1) / KS;
  const int kb0 = kh * kper, kb1 = min(kblocks, kb0 + kper);
  float acc[MT][4];
#pragma unroll
  for (int m = 0; m < MT; ++m) acc[m][0] = acc[m][1] = acc[m][2] = acc[m][3] = 0.f;
  const uint4* wp = reinterpret_cast<const uint4*>(W + (long long)(n0 + g) * K + 8 * t4);
  const bool col_ok = n0 + g < N;
  // Issue a whole batch of weight loads before touching them: the op is one DRAM round trip, so
  // what matters is how many bytes are in flight, not how fast they are consumed. The first batch
  // is requested BEFORE the grid dependency resolves (PDL): the weights are immutable, so their DRAM
  // latency hides behind the producer kernel's tail; A / residual are only touched after pdl_wait().
  constexpr int PF = MT <= 4 ? 12 : (MT <= 8 ? 8 : 4);
  uint4 wv[PF];
#pragma unroll
  for (int i = 0; i < PF; ++i) {
    wv[i] = make_uint4(0, 0, 0, 0);
    if (col_ok && kb0 + i < kb1) wv[i] = __ldg(wp + (kb0 + i) * 4);  // 32 k = 64 B = 4 uint4 per row
  }
  pdl_launch_dependents();
  pdl_wait();
#pragma unroll 1
  for (int kbase = kb0; kbase < kb1; kbase += PF) {
#pragma unroll
    for (int i = 0; i < PF; ++i) {
      const int kb = kbase + i;
      if (kb < kb1) {
#pragma unroll
        for (int m = 0; m < MT; ++m) {
          const int r0 = 16 * m + g, r1 = r0 + 8;
          uint4 a_lo = make_uint4(0, 0, 0, 0), a_hi = make_uint4(0, 0, 0, 0);
          if (r0 < rows) a_lo = *reinterpret_cast<const uint4*>(A + (long long)r0 * lda + kb * 32 + 8 * t4);
          if (r1 < rows) a_hi = *reinterpret_cast<const uint4*>(A + (long long)r1 * lda + kb * 32 + 8 * t4);
          mma_bf16_16816(acc[m], a_lo.x, a_hi.x, a_lo.y, a_hi.y, wv[i].x, wv[i].y);
          mma_bf16_16816(acc[m], a_lo.z, a_hi.z, a_lo.w, a_hi.w, wv[i].z, wv[i].w);
        }
      }
    }
    if (kbase + PF < kb1) {
#pragma unroll
      for (int i = 0; i < PF; ++i) {
        wv[i] = make_uint4(0, 0, 0, 0);
        if (col_ok && kbase + PF + i < kb1) wv[i] = __ldg(wp + (kbase + PF + i) * 4);
      }
    }
  }
  if (kh > 0) {
#pragma unroll
    for (int m = 0; m < MT; ++m)
      *reinterpret_cast<float4*>(red[kh - 1][ct][m][lane]) = make_float4(acc[m][0], acc[m][1], acc[m][2], acc[m][3]);
  }
  __syncthreads();
  if (kh > 0) return;
#pragma unroll
  for (int s = 0; s < KS - 1; ++s)      // fixed summation order: deterministic
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      const float4 r4 = *reinterpret_cast<const float4*>(red[s][ct][m][lane]);
      acc[m][0] += r4.x; acc[m][1] += r4.y; acc[m][2] += r4.z; acc[m][3] += r4.w;
    }
  const int col = n0 + 2 * t4;
  if (col >= N) return;
  float b0 = 0.f, b1 = 0.f;
  if (bias) { b0 = __ldg(bias + col); b1 = __ldg(bias + col + 1); }
#pragma unroll
  for (int m = 0; m < MT; ++m) {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int r = 16 * m + g + 8 * half;
      if (r >= rows) continue;
      float v0 = acc[m][2 * half] + b0;
      float v1 = acc[m][2 * half + 1] + b1;
      if (act == ACT_GELU) { v0 = gelu_erf(v0); v1 = gelu_erf(v1); }
      const long long o = (long long)r * N + col;
      if (residual) {
        const float2 rr = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(residual + o));
        v0 += rr.x; v1 += rr.y;
      }
      if (out_f32) *reinterpret_cast<float2*>(static_cast<float*>(out) + o) = make_float2(v0, v1);
      else *reinterpret_cast<__nv_bfloat162*>(static_cast<__nv_bfloat16*>(out) + o) = __floats2bfloat162_rn(v0, v1);
    }
  }
}

int dec_linear(const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, const float* bias,
               const __nv_bfloat16* residual, void* out, int rows, int N, int K, int act, int out_f32,
               cudaStream_t stream, bool pdl) {
  JANUS_CHECK_ARG(K % 32 == 0 && N % 2 == 0 && lda % 8 == 0, "dec_linear: K %% 32, N %% 2, lda %% 8 required");
  JANUS_CHECK_ARG(rows > 0 && rows <= 256, "dec_linear: 1..256 rows supported (got %d)", rows);
  const int grid = (N + DL_COLS - 1) / DL_COLS;
  const int mt = (rows + 15) / 16;
#define DL_LAUNCH(MT, KS)                                                                                          \
  do {                                                                                                             \
    if (pdl) {                                                                                                     \
      JANUS_CUDA(launch_pdl(dec_linear_kernel<MT, KS>, dim3(grid), dim3(2 * KS * 32), 0, stream, A, lda, W, bias,    \
                            residual, out, rows, N, K, act, out_f32));                                             \
    } else {                                                                                                       \
      dec_linear_kernel<MT, KS><<<grid, 2 * KS * 32, 0, stream>>>(A, lda, W, bias, residual, out, rows, N, K, act,  \
                                                                  out_f32);                                        \
    }                                                                                                              \
  } while (0)
  if (K > 1024 && mt <= 4) {   // long K (MLP down-projection): 8-way K split
    if (mt <= 1) DL_LAUNCH(1, 8);
    else if (mt <= 2) DL_LAUNCH(2, 8);
    else DL_LAUNCH(4, 8);
  } else if (mt <= 1) DL_LAUNCH(1, 2);
  else if (mt <= 2) DL_LAUNCH(2, 2);
  else if (mt <= 4) DL_LAUNCH(4, 2);
  else if (mt <= 8) DL_LAUNCH(8, 2);
  else DL_LAUNCH(16, 2);
#undef DL_LAUNCH
  JANUS_LAUNCH_CHECK("dec_linear_kernel");
  return JANUS_OK;
}

}  // namespace janus

extern "C" int32_t janus_dec_linear_bf16(const void* A, int32_t lda, const void* W, const float* bias,
                                         const void* residual, void* out, int32_t rows, int32_t N, int32_t K,
                                         int32_t act, int32_t out_f32, void* stream) {
  JANUS_CHECK_ARG(A && W && out, "dec_linear: null pointer");
  return janus::dec_linear(static_cast<const __nv_bfloat16*>(A), lda, static_cast<const __nv_bfloat16*>(W), bias,
                           static_cast<const __nv_bfloat16*>(residual), out, rows, N, K, act, out_f32,
                           (cudaStream_t)stream, true);
}
