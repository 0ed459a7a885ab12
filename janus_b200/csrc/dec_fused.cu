// K3, latency path: fused "segment" kernel for the decode step.
//
// One greedy decode step of Whisper (CTranslate2 `Whisper.generate` in the reference's stack,
// /root/reference/backend/services/transcriber.py:53-64) is ~140 tiny dependent ops for a batch of
// streams. Launched one kernel per op, the step is bound by the dependent-launch latency (3.4 us per
// kernel inside a CUDA graph on B200, measured: scripts/gpu_pdl_probe.py) and by every op paying its own
// ramp-up, not by any roofline. Only the cross-attention over the 1500 encoder states is HBM-bound and
// stays its own kernel (decoder.cu). Everything between two cross-attentions runs here as ONE
// persistent kernel that executes a short program of phases separated by a grid-wide barrier (~0.5 us):
//
//   co-proj(+res) | LN2+fc1+GELU | fc2(+res) | LN1+QKV | self-attention | o-proj(+res) | LNc+q-proj
//
// Linear phases: the CTA stages a 64-row x d slab of the activations in shared memory with cp.async
// (L2 -> smem, bypassing L1), applies the LayerNorm in place where the op has one (each CTA owns complete
// rows, so no separate LN pass / buffer / barrier exists), and feeds mma.sync.m16n8k16 from it with
// conflict-free 16-byte LDS. Weights stream HBM -> registers exactly once with 16-byte loads (K permuted
// identically for A and W, as in dec_linear.cu); the first weight lines of the NEXT phase are
// prefetched into L2 before the barrier so their DRAM latency hides behind it.
// Summation order is fixed (no atomics): results are deterministic and independent of the grid size.
#include "kernels.cuh"

namespace janus {

constexpr int FZ_THREADS = 256;           // 8 warps
constexpr int FZ_ROWS = 64;               // activation rows per tile (4 m16 tiles)
constexpr int FZ_PAD = 32;                // bf16 elements of row padding: stride = (d + 32) * 2 B -> conflict-free LDS.128
constexpr int FZ_PF = 12;                 // weight k-blocks (of 32) in flight per thread
constexpr int FZ_DH = 64;

__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, bool valid) {
  const int sz = valid ? 16 : 0;   // src-size 0 -> zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];\n" ::"l"(p)); }
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];\n" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// Grid-wide barrier on a monotonically increasing 64-bit counter (never reset: the kernel derives its
// base from the value it finds, which is a multiple of gridDim.x whenever no launch is in flight).
struct GridBarrier {
  unsigned long long* ctr;
  unsigned long long target;
  __device__ void init(unsigned long long* c) {
    ctr = c;
    target = 0;
    if (threadIdx.x == 0) target = (ld_acquire_u64(c) / gridDim.x) * gridDim.x;  // read BEFORE our first arrival
  }
  __device__ void sync() {
    __syncthreads();
    if (threadIdx.x == 0) {
      target += gridDim.x;
      __threadfence();                    // release our CTA's writes
      atomicAdd(ctr, 1ULL);
      while (ld_acquire_u64(ctr) < target) { __nanosleep(20); }
      __threadfence();                    // gpu-scope fence also invalidates this SM's L1
    }
    __syncthreads();
  }
};

// ------------------------------------------------------------------------------------------
// linear phase
// ------------------------------------------------------------------------------------------
// Warp roles for a tile of 64 rows x NB columns (cfg picks NB):
//   cfg 0: NB = 32   4 n8-tiles x 2 K halves          (N = d outputs: 96 tiles at 256 rows)
//   cfg 1: NB = 64   8 n8-tiles, full K per warp      (QKV)
//   cfg 2: NB = 128  16 n8-tiles, 2 per warp, full K  (fc1)
__device__ __forceinline__ int cfg_nb(int cfg) { return cfg == 0 ? 32 : (cfg == 1 ? 64 : 128); }

__device__ __forceinline__ void linear_prefetch(const FusedLinear& L, int rows) {
  // first tile of this CTA in phase L: touch the first weight lines of every warp's rows (L2 prefetch)
  const int nb = cfg_nb(L.cfg);
  const int n_rb = (rows + FZ_ROWS - 1) / FZ_ROWS;
  const int n_tiles = n_rb * (L.N / nb);
  if ((int)blockIdx.x >= n_tiles) return;
  const int cb = blockIdx.x / n_rb;
  const int kc = L.K < L.kc ? L.K : L.kc;
  const __nv_bfloat16* W = static_cast<const __nv_bfloat16*>(L.W) + (long long)cb * nb * L.K;
  const int lines_per_row = (kc * 2 + 127) / 128;
  const int total = nb * lines_per_row;
  for (int i = threadIdx.x; i < total; i += FZ_THREADS) {
    const int r = i / lines_per_row, ln = i - r * lines_per_row;
    prefetch_l2(W + (long long)r * L.K + ln * 64);
  }
}

template <int NT, int KS>
__device__ __forceinline__ void linear_tile(const FusedLinear& L, int rows, int row0, int n0, int d_ln,
                                            __nv_bfloat16* sA, float* sRed) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  constexpr int WPK = 8 / KS;                 // warps per K split
  const int wn = warp % WPK, kh = warp / WPK;
  const int kc = L.K < L.kc ? L.K : L.kc;     // K chunk staged in smem
  const int stride = kc + FZ_PAD;             // smem row stride (elements)
  const __nv_bfloat16* A = static_cast<const __nv_bfloat16*>(L.A);
  const __nv_bfloat16* W = static_cast<const __nv_bfloat16*>(L.W);
  float acc[4][NT][4];
#pragma unroll
  for (int m = 0; m < 4; ++m)
#pragma unroll
    for (int j = 0; j < NT; ++j) acc[m][j][0] = acc[m][j][1] = acc[m][j][2] = acc[m][j][3] = 0.f;

  const int kblocks = kc >> 5;
  const int kper = (kblocks + KS - 1) / KS;
  const int kb0 = kh * kper, kb1 = min(kblocks, kb0 + kper);

  for (int k0 = 0; k0 < L.K; k0 += kc) {
    __syncthreads();   // previous readers of sA are done
    {   // stage A[row0 .. row0+63][k0 .. k0+kc) -> smem (zero rows past `rows`)
      const int chunks_per_row = kc >> 3;     // 16-byte chunks
      const int total = FZ_ROWS * chunks_per_row;
      for (int i = threadIdx.x; i < total; i += FZ_THREADS) {
        const int r = i / chunks_per_row, c = i - r * chunks_per_row;
        const bool ok = row0 + r < rows;
        const __nv_bfloat16* src = A + (long long)(ok ? row0 + r : 0) * L.lda + k0 + c * 8;
        cp_async16(sA + r * stride + c * 8, src, ok);
      }
      cp_async_wait_all();
    }
    __syncthreads();
    if (L.ln_w != nullptr) {
      // LayerNorm in place (fp32 statistics, eps 1e-5; same arithmetic as layernorm_bf16_kernel): warp w
      // owns rows 8w .. 8w+7; requires kc == K == d
      const int npairs = d_ln >> 6;
      const float2* w2 = reinterpret_cast<const float2*>(L.ln_w);
      const float2* b2 = reinterpret_cast<const float2*>(L.ln_b);
      for (int rr = 0; rr < 8; ++rr) {
        __nv_bfloat162* xr = reinterpret_cast<__nv_bfloat162*>(sA + (warp * 8 + rr) * stride);
        float2 v[20];
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < 20; ++i)
          if (i < npairs) { v[i] = __bfloat1622float2(xr[lane + 32 * i]); sum += v[i].x + v[i].y; }
        const float mean = warp_sum(sum) / (float)d_ln;
        float sq = 0.f;
#pragma unroll
        for (int i = 0; i < 20; ++i)
          if (i < npairs) { const float a = v[i].x - mean, c = v[i].y - mean; sq += a * a + c * c; }
        const float rstd = rsqrtf(warp_sum(sq) / (float)d_ln + 1e-5f);
#pragma unroll
        for (int i = 0; i < 20; ++i)
          if (i < npairs) {
            const float2 ww = __ldg(w2 + lane + 32 * i), bb = __ldg(b2 + lane + 32 * i);
            xr[lane + 32 * i] = __floats2bfloat162_rn((v[i].x - mean) * rstd * ww.x + bb.x,
                                                      (v[i].y - mean) * rstd * ww.y + bb.y);
          }
      }
      __syncthreads();
    }
    // ---- mainloop: weights HBM/L2 -> registers, A fragments from smem
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int ncol = n0 + (wn + j * WPK) * 8 + g;   // this thread's weight row
      const uint4* wp = reinterpret_cast<const uint4*>(W + (long long)ncol * L.K + k0 + 8 * t4);
#pragma unroll 1
      for (int kbase = kb0; kbase < kb1; kbase += FZ_PF) {
        uint4 wv[FZ_PF];
#pragma unroll
        for (int i = 0; i < FZ_PF; ++i) {
          wv[i] = make_uint4(0, 0, 0, 0);
          if (kbase + i < kb1) wv[i] = __ldg(wp + (kbase + i) * 4);
        }
#pragma unroll
        for (int i = 0; i < FZ_PF; ++i) {
          const int kb = kbase + i;
          if (kb < kb1) {
#pragma unroll
            for (int m = 0; m < 4; ++m) {
              const __nv_bfloat16* ap = sA + (16 * m + g) * stride + kb * 32 + 8 * t4;
              const uint4 a_lo = *reinterpret_cast<const uint4*>(ap);
              const uint4 a_hi = *reinterpret_cast<const uint4*>(ap + 8 * stride);
              mma16816(acc[m][j], a_lo.x, a_hi.x, a_lo.y, a_hi.y, wv[i].x, wv[i].y);
              mma16816(acc[m][j], a_lo.z, a_hi.z, a_lo.w, a_hi.w, wv[i].z, wv[i].w);
            }
          }
        }
      }
    }
  }
  // ---- combine the K splits (fixed order) and run the epilogue from registers
  if (KS > 1) {
    __syncthreads();
    if (kh > 0) {
#pragma unroll
      for (int m = 0; m < 4; ++m)
#pragma unroll
        for (int j = 0; j < NT; ++j)
          *reinterpret_cast<float4*>(sRed + ((((kh - 1) * WPK + wn) * 4 + m) * NT + j) * 128 + lane * 4) =
              make_float4(acc[m][j][0], acc[m][j][1], acc[m][j][2], acc[m][j][3]);
    }
    __syncthreads();
    if (kh > 0) return;
#pragma unroll
    for (int s = 0; s < KS - 1; ++s)
#pragma unroll
      for (int m = 0; m < 4; ++m)
#pragma unroll
        for (int j = 0; j < NT; ++j) {
          const float4 r4 = *reinterpret_cast<const float4*>(sRed + (((s * WPK + wn) * 4 + m) * NT + j) * 128 + lane * 4);
          acc[m][j][0] += r4.x; acc[m][j][1] += r4.y; acc[m][j][2] += r4.z; acc[m][j][3] += r4.w;
        }
  }
  const __nv_bfloat16* res = static_cast<const __nv_bfloat16*>(L.residual);
  __nv_bfloat16* out = static_cast<__nv_bfloat16*>(L.out);
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const int col = n0 + (wn + j * WPK) * 8 + 2 * t4;
    float b0 = 0.f, b1 = 0.f;
    if (L.bias) { b0 = __ldg(L.bias + col); b1 = __ldg(L.bias + col + 1); }
#pragma unroll
    for (int m = 0; m < 4; ++m)
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int r = row0 + 16 * m + g + 8 * half;
        if (r >= rows) continue;
        float v0 = acc[m][j][2 * half] + b0, v1 = acc[m][j][2 * half + 1] + b1;
        if (L.act == ACT_GELU) { v0 = gelu_erf(v0); v1 = gelu_erf(v1); }
        const long long o = (long long)r * L.N + col;
        if (res) {
          const uint32_t u = __ldcg(reinterpret_cast<const uint32_t*>(res + o));
          const float2 rr = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u));
          v0 += rr.x; v1 += rr.y;
        }
        *reinterpret_cast<__nv_bfloat162*>(out + o) = __floats2bfloat162_rn(v0, v1);
      }
  }
}

__device__ __forceinline__ void linear_phase(const FusedLinear& L, int rows, int d_ln, __nv_bfloat16* sA,
                                             float* sRed) {
  const int nb = cfg_nb(L.cfg);
  const int n_rb = (rows + FZ_ROWS - 1) / FZ_ROWS;
  const int n_tiles = n_rb * (L.N / nb);
  for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const int rb = t % n_rb, cb = t / n_rb;   // consecutive CTAs share the weight columns (L2 reuse)
    if (L.cfg == 0) linear_tile<1, 2>(L, rows, rb * FZ_ROWS, cb * nb, d_ln, sA, sRed);
    else if (L.cfg == 1) linear_tile<1, 1>(L, rows, rb * FZ_ROWS, cb * nb, d_ln, sA, sRed);
    else linear_tile<2, 1>(L, rows, rb * FZ_ROWS, cb * nb, d_ln, sA, sRed);
  }
}

// ------------------------------------------------------------------------------------------
// self-attention phase: one warp per (stream, head); same arithmetic as dec_self_attn_kernel
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void self_attn_phase(const FusedSelfAttn& S, int rows, int d, int H, int pos, float* smem_f) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* ps = smem_f + warp * (512 + FZ_DH);
  float* qs = ps + 512;
  const __nv_bfloat16* qkv = static_cast<const __nv_bfloat16*>(S.qkv);
  __nv_bfloat16* out = static_cast<__nv_bfloat16*>(S.out);
  const int n_tasks = rows * H;
  for (int task = blockIdx.x * 8 + warp; task < n_tasks; task += gridDim.x * 8) {
    const int b = task / H, h = task - b * H;
    const __nv_bfloat16* row = qkv + (long long)b * 3 * d + h * FZ_DH;
    __nv_bfloat16* K = static_cast<__nv_bfloat16*>(S.k_cache) + ((long long)b * H + h) * S.n_ctx * FZ_DH;
    __nv_bfloat16* V = static_cast<__nv_bfloat16*>(S.v_cache) + ((long long)b * H + h) * S.n_ctx * FZ_DH;
    const uint32_t kq = __ldcg(reinterpret_cast<const uint32_t*>(row + d) + lane);
    const uint32_t vq = __ldcg(reinterpret_cast<const uint32_t*>(row + 2 * d) + lane);
    reinterpret_cast<uint32_t*>(K + (long long)pos * FZ_DH)[lane] = kq;
    reinterpret_cast<uint32_t*>(V + (long long)pos * FZ_DH)[lane] = vq;
    {
      const uint32_t u = __ldcg(reinterpret_cast<const uint32_t*>(row) + lane);
      const float2 qq = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u));
      qs[2 * lane] = qq.x * 0.125f;
      qs[2 * lane + 1] = qq.y * 0.125f;
    }
    __syncwarp();
    float mx = -INFINITY;
    for (int j = lane; j <= pos; j += 32) {
      float s = 0.f;
      const uint4* kr = reinterpret_cast<const uint4*>(K + (long long)j * FZ_DH);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const uint4 u = __ldcg(kr + c);
        const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = __bfloat1622float2(hh[e]);
          s = fmaf(f.x, qs[c * 8 + 2 * e], s);
          s = fmaf(f.y, qs[c * 8 + 2 * e + 1], s);
        }
      }
      ps[j] = s;
      mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j <= pos; j += 32) {
      const float p = __expf(ps[j] - mx);
      ps[j] = p;
      sum += p;
    }
    sum = warp_sum(sum);
    __syncwarp();
    float o0 = 0.f, o1 = 0.f;
    for (int j = 0; j <= pos; ++j) {
      const uint32_t u = __ldcg(reinterpret_cast<const uint32_t*>(V + (long long)j * FZ_DH) + lane);
      const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u));
      const float p = ps[j];
      o0 = fmaf(p, f.x, o0);
      o1 = fmaf(p, f.y, o1);
    }
    const float inv = 1.f / sum;
    reinterpret_cast<__nv_bfloat162*>(out + (long long)b * d + h * FZ_DH)[lane] = __floats2bfloat162_rn(o0 * inv, o1 * inv);
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------------------
// plain LayerNorm phase (final ln before the vocabulary projection): warp per row, global -> global
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void layernorm_phase(const FusedLayerNorm& P, int rows, int d) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int npairs = d >> 6;
  const float2* w2 = reinterpret_cast<const float2*>(P.w);
  const float2* b2 = reinterpret_cast<const float2*>(P.b);
  for (int row = blockIdx.x * 8 + warp; row < rows; row += gridDim.x * 8) {
    const uint32_t* xr = reinterpret_cast<const uint32_t*>(static_cast<const __nv_bfloat16*>(P.x) + (long long)row * d);
    float2 v[20];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < 20; ++i)
      if (i < npairs) {
        const uint32_t u = __ldcg(xr + lane + 32 * i);
        v[i] = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u));
        sum += v[i].x + v[i].y;
      }
    const float mean = warp_sum(sum) / (float)d;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < 20; ++i)
      if (i < npairs) { const float a = v[i].x - mean, c = v[i].y - mean; sq += a * a + c * c; }
    const float rstd = rsqrtf(warp_sum(sq) / (float)d + 1e-5f);
    __nv_bfloat162* yr = reinterpret_cast<__nv_bfloat162*>(static_cast<__nv_bfloat16*>(P.y) + (long long)row * d);
#pragma unroll
    for (int i = 0; i < 20; ++i)
      if (i < npairs) {
        const float2 ww = __ldg(w2 + lane + 32 * i), bb = __ldg(b2 + lane + 32 * i);
        yr[lane + 32 * i] = __floats2bfloat162_rn((v[i].x - mean) * rstd * ww.x + bb.x,
                                                  (v[i].y - mean) * rstd * ww.y + bb.y);
      }
  }
}

// ------------------------------------------------------------------------------------------
// the segment kernel
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(FZ_THREADS, 1) dec_fused_segment_kernel(const __grid_constant__ FusedProgram prog) {
  extern __shared__ __align__(16) uint8_t fz_smem[];
  __nv_bfloat16* sA = reinterpret_cast<__nv_bfloat16*>(fz_smem);
  float* sRed = reinterpret_cast<float*>(fz_smem + (size_t)FZ_ROWS * (prog.d + FZ_PAD) * 2);
  GridBarrier bar;
  bar.init(prog.barrier);
  if (prog.ph[0].type == FZ_LINEAR) linear_prefetch(prog.ph[0].lin, prog.rows);
  pdl_wait();   // (launched with PDL behind the cross-attention kernel: weights are requested early)
  const int pos = *prog.pos;
#pragma unroll 1
  for (int p = 0; p < prog.n_phases; ++p) {
    const FusedPhase& ph = prog.ph[p];
    if (p > 0) bar.sync();
    if (ph.type == FZ_LINEAR) linear_phase(ph.lin, prog.rows, prog.d, sA, sRed);
    else if (ph.type == FZ_SELF_ATTN) self_attn_phase(ph.sa, prog.rows, prog.d, prog.n_heads, pos, reinterpret_cast<float*>(fz_smem));
    else layernorm_phase(ph.ln, prog.rows, prog.d);
    if (p + 1 < prog.n_phases && prog.ph[p + 1].type == FZ_LINEAR) linear_prefetch(prog.ph[p + 1].lin, prog.rows);
  }
}

size_t dec_fused_smem_bytes(int d) {
  const size_t a = (size_t)FZ_ROWS * (d + FZ_PAD) * 2;
  const size_t red = 4 /*warps kh=1*/ * 4 /*m*/ * 128 * sizeof(float);   // cfg 0 only (KS = 2, NT = 1)
  const size_t sa = 8 * (512 + FZ_DH) * sizeof(float);                    // self-attention scratch aliases sA
  return (a + red > sa ? a + red : sa);
}

int dec_fused_segment(const FusedProgram& prog, cudaStream_t stream) {
  JANUS_CHECK_ARG(prog.n_phases > 0 && prog.n_phases <= FZ_MAX_PHASES, "fused segment: bad phase count %d", prog.n_phases);
  JANUS_CHECK_ARG(prog.d % 64 == 0 && prog.d <= 1280, "fused segment: d=%d must be a multiple of 64 and <= 1280", prog.d);
  for (int i = 0; i < prog.n_phases; ++i) {
    const FusedPhase& ph = prog.ph[i];
    if (ph.type != FZ_LINEAR) continue;
    const FusedLinear& L = ph.lin;
    const int nb = L.cfg == 0 ? 32 : (L.cfg == 1 ? 64 : 128);
    JANUS_CHECK_ARG(L.N % nb == 0 && L.kc == prog.d && L.K % L.kc == 0 && L.lda % 8 == 0,
                    "fused segment: linear phase %d has unsupported shape N=%d K=%d", i, L.N, L.K);
    JANUS_CHECK_ARG(!L.ln_w || L.K == prog.d, "fused segment: fused LayerNorm needs K == d");
  }
  const size_t smem = dec_fused_smem_bytes(prog.d);
  static size_t attr_smem = 0;
  if (smem > attr_smem) {
    JANUS_CUDA(cudaFuncSetAttribute(dec_fused_segment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_smem = smem;
  }
  // all CTAs must be co-resident (grid barrier): one per SM, 256 threads, ~110 KB of shared memory
  dec_fused_segment_kernel<<<num_sms(), FZ_THREADS, smem, stream>>>(prog);
  JANUS_LAUNCH_CHECK("dec_fused_segment_kernel");
  return JANUS_OK;
}

}  // namespace janus

// Diagnostic (not on the product path): a segment of `n_phases` identical linear phases, for timing one
// phase + barrier in isolation (scripts/gpu_fused_probe.py).
extern "C" int32_t janus_debug_fused_linear(const void* A, const void* W, const float* bias, void* out,
                                            const float* ln_w, const float* ln_b, int32_t rows, int32_t N, int32_t K,
                                            int32_t d, int32_t cfg, int32_t n_phases, void* barrier, const int32_t* pos,
                                            void* stream) {
  using namespace janus;
  FusedProgram P = {};
  P.rows = rows; P.d = d; P.n_heads = d / 64; P.pos = pos;
  P.barrier = static_cast<unsigned long long*>(barrier);
  for (int i = 0; i < n_phases && i < FZ_MAX_PHASES; ++i) {
    FusedPhase ph = {};
    ph.type = FZ_LINEAR;
    ph.lin.A = A; ph.lin.lda = K; ph.lin.ln_w = ln_w; ph.lin.ln_b = ln_b; ph.lin.W = W; ph.lin.bias = bias;
    ph.lin.out = out; ph.lin.N = N; ph.lin.K = K; ph.lin.kc = d; ph.lin.cfg = cfg;
    P.ph[P.n_phases++] = ph;
  }
  return dec_fused_segment(P, (cudaStream_t)stream);
}
