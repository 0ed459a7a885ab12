// K3: Whisper KV-cache decoder (greedy, temperature 0) for a batch of independent windows.
//
// Replaces CTranslate2 `Whisper.generate` + faster-whisper's decode-time logit rules, which the
// reference runs lazily while iterating `segments` (/root/reference/backend/services/transcriber.py:
// 53-64, beam_size=1): token + learned positional embedding, n_layers x [LN -> causal self-attention
// over the KV cache, LN -> cross-attention over the 1500 encoder states, LN -> MLP], LN, logits =
// x E^T, SuppressBlank / SuppressTokens / ApplyTimestampRules, log-softmax, argmax.
//
// The step is memory-bound. Per stream and step it must read the cross K/V of every layer
// (2 * L * 1500 * d bf16 = 55 MB for Whisper-small), so the cross-attention kernel owns the layout:
//   cross K is stored TRANSPOSED  kT[l][b][h][64][1536]  -> scores: lanes <-> keys, 16-byte loads of 8
//                                                            consecutive keys per dim, no shuffles
//   cross V is stored row-major   v [l][b][h][1500][64]  -> PV: 8 lanes cover one 128-byte row
// Both are written directly in those layouts by the tcgen05 GEMM epilogue (STORE_CROSS_KV).
// The skinny (M = batch) linear layers use a dedicated register-streaming kernel (dec_linear.cu): every
// weight byte is read exactly once per step by hundreds of warps; logits are produced in fp32 and the
// rule filter + log-softmax + argmax run in one pass of one CTA per stream.
// All per-step scalars (position, done flags) live in device memory so the whole step is one
// CUDA graph that the host replays.
#include <stdlib.h>

#include "kernels.cuh"

namespace janus {

constexpr int DH = 64;

// ------------------------------------------------------------------------------------------
// embedding
// ------------------------------------------------------------------------------------------
// (mutable operands of the PDL chain -- tokens, pos, activations, KV caches -- carry no __restrict__ so
// they are never read through the non-coherent path; see common.cuh)
__global__ void dec_embed_kernel(const int32_t* tokens, int max_len, const int32_t* pos_p,
                                 const __nv_bfloat16* __restrict__ emb, const float* __restrict__ pos_emb, int d,
                                 __nv_bfloat16* x) {
  const int b = blockIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const int pos = *pos_p;
  const int tok = tokens[(long long)b * max_len + pos];
  const __nv_bfloat16* e = emb + (long long)tok * d;
  const float* pe = pos_emb + (long long)pos * d;
  for (int i = threadIdx.x; i < d; i += blockDim.x)
    x[(long long)b * d + i] = __float2bfloat16_rn(__bfloat162float(e[i]) + pe[i]);
}

// ------------------------------------------------------------------------------------------
// causal self-attention over the KV cache: one warp per (stream, head), 4 per CTA
// ------------------------------------------------------------------------------------------
// At 1024 streams this is 12288 independent (stream, head) problems of <= 131 keys each per layer: a
// streaming kernel (~200 MB of cache per launch), so what matters is resident warps and loads in flight:
// scores with one key row (128 B) per lane, PV with 4 rows x 16 B per lane-group and 4-way unrolling.
constexpr int SA_WARPS = 4;
constexpr int SA_MAXCTX = 512;

// BEAM: the rows of a stream's beams (row = stream * beam + k) keep appending to their own cache slot; which slot
// holds position j of row's hypothesis is looked up in indir[row][j] (beam index within the stream), so a beam
// reorder rewrites `beam * n_ctx` bytes of indices per stream instead of copying K/V rows (dec_beam.cu).
template <bool BEAM>
__global__ void __launch_bounds__(SA_WARPS * 32)
dec_self_attn_kernel(const __nv_bfloat16* qkv, __nv_bfloat16* kc, __nv_bfloat16* vc, const int32_t* pos_p, int d,
                     int n_ctx, int H, int n_items, __nv_bfloat16* out, const uint8_t* indir, int beam) {
  __shared__ float ps_all[SA_WARPS][SA_MAXCTX];
  __shared__ float qs_all[SA_WARPS][DH];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  pdl_wait();
  const int item = blockIdx.x * SA_WARPS + warp;
  if (item >= n_items) return;
  float* ps = ps_all[warp];
  float* qs = qs_all[warp];
  const int b = item / H, h = item - b * H;
  const int pos = *pos_p;
  const __nv_bfloat16* row = qkv + (long long)b * 3 * d + h * DH;
  __nv_bfloat16* K = kc + ((long long)b * H + h) * n_ctx * DH;
  __nv_bfloat16* V = vc + ((long long)b * H + h) * n_ctx * DH;
  // append this step's k, v
  reinterpret_cast<__nv_bfloat162*>(K + (long long)pos * DH)[lane] = reinterpret_cast<const __nv_bfloat162*>(row + d)[lane];
  reinterpret_cast<__nv_bfloat162*>(V + (long long)pos * DH)[lane] = reinterpret_cast<const __nv_bfloat162*>(row + 2 * d)[lane];
  {
    const float2 qq = __bfloat1622float2(reinterpret_cast<const __nv_bfloat162*>(row)[lane]);
    qs[2 * lane] = qq.x * 0.125f;
    qs[2 * lane + 1] = qq.y * 0.125f;
  }
  __syncwarp();
  // scores: 8 lanes share one 128-byte key row (fully coalesced: 4 rows per load instruction), partial dot
  // products over 8 dims each, 3 shuffles per row. (One row per lane costs 32 L1 wavefronts per 16-byte
  // load and capped the kernel at half the HBM rate.)
  const int jg = lane >> 3, c8 = lane & 7;
  float qf[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) qf[i] = qs[c8 * 8 + i];
  const uint4* kp = reinterpret_cast<const uint4*>(K) + c8;
  // BEAM: uint4 offset of slot s relative to this row's own cache
  const long long slot_stride = (long long)H * n_ctx * 8;
  const uint8_t* ind = BEAM ? indir + (long long)b * n_ctx : nullptr;
  const int self_slot = BEAM ? b % beam : 0;
#pragma unroll 4
  for (int j0 = 0; j0 <= pos; j0 += 4) {   // warp-uniform trip count (the shuffles below need every lane)
    const int j = j0 + jg;
    uint4 u = make_uint4(0, 0, 0, 0);
    if (j <= pos) {
      if (BEAM) u = kp[(long long)j * 8 + ((int)ind[j] - self_slot) * slot_stride];
      else u = kp[(long long)j * 8];
    }
    const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
    float s = 0.f;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(hh[e]);
      s = fmaf(f.x, qf[2 * e], s);
      s = fmaf(f.y, qf[2 * e + 1], s);
    }
    s += __shfl_xor_sync(FULL_MASK, s, 1);
    s += __shfl_xor_sync(FULL_MASK, s, 2);
    s += __shfl_xor_sync(FULL_MASK, s, 4);
    if (c8 == 0 && j <= pos) ps[j] = s;
  }
  __syncwarp();
  float mx = -INFINITY;
  for (int j = lane; j <= pos; j += 32) mx = fmaxf(mx, ps[j]);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int j = lane; j <= pos; j += 32) {
    const float p = __expf(ps[j] - mx);
    ps[j] = p;
    sum += p;
  }
  sum = warp_sum(sum);
  __syncwarp();
  // PV: lane group jg = lane / 8 takes rows jg, jg + 4, ...; lane c = lane % 8 owns dims 8c .. 8c+7
  float o[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i] = 0.f;
  const uint4* vp = reinterpret_cast<const uint4*>(V) + c8;
#pragma unroll 4
  for (int j = jg; j <= pos; j += 4) {
    const uint4 u = BEAM ? vp[(long long)j * 8 + ((int)ind[j] - self_slot) * slot_stride] : vp[(long long)j * 8];
    const float p = ps[j];
    const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(hh[e]);
      o[2 * e] = fmaf(p, f.x, o[2 * e]);
      o[2 * e + 1] = fmaf(p, f.y, o[2 * e + 1]);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    o[i] += __shfl_xor_sync(FULL_MASK, o[i], 8);
    o[i] += __shfl_xor_sync(FULL_MASK, o[i], 16);
  }
  if (jg == 0) {
    const float inv = 1.f / sum;
    uint4 w;
    __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&w);
#pragma unroll
    for (int e = 0; e < 4; ++e) hh[e] = __floats2bfloat162_rn(o[2 * e] * inv, o[2 * e + 1] * inv);
    reinterpret_cast<uint4*>(out + (long long)b * d + h * DH)[c8] = w;
  }
}

// ------------------------------------------------------------------------------------------
// causal self-attention for beam search: one warp per (WINDOW, head) serves all K beams. The beams of a window
// descend from a common ancestor: for positions j < c (c = first position where their indirection entries differ;
// if all beams use the same slot at j they share everything before j) a K / V row is loaded ONCE and used by all K
// queries; only the last pos + 1 - c positions are read per beam. dec_self_attn_kernel<true> re-reads the shared
// prefix once per beam (the duplicates hit L2, but at position ~115 it still costs 12 x 0.158 ms of an 8.6 ms step
// for 2560 rows of Whisper-small).
// ------------------------------------------------------------------------------------------
constexpr int SB_WARPS = 4;

template <int K>
__global__ void __launch_bounds__(SB_WARPS * 32)
dec_self_attn_beam_kernel(const __nv_bfloat16* qkv, __nv_bfloat16* kc, __nv_bfloat16* vc, const int32_t* pos_p, int d,
                          int n_ctx, int H, int n_items, __nv_bfloat16* out, const uint8_t* indir) {
  extern __shared__ float sb_smem[];   // per warp: ps[K][n_ctx], qs[K][64]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  pdl_wait();
  const int item = blockIdx.x * SB_WARPS + warp;
  if (item >= n_items) return;
  float* ps = sb_smem + (size_t)warp * K * (n_ctx + DH);
  float* qs = ps + K * n_ctx;
  const int w = item / H, h = item - w * H;
  const long long row0 = (long long)w * K;
  const int pos = *pos_p;
  const long long slot_stride = (long long)H * n_ctx * DH;                  // elements between two rows' caches
  __nv_bfloat16* K0 = kc + (row0 * H + h) * (long long)n_ctx * DH;           // slot 0 of this window, this head
  __nv_bfloat16* V0 = vc + (row0 * H + h) * (long long)n_ctx * DH;
  // append this step's k, v of every beam to its own slot; stage the scaled queries
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const __nv_bfloat16* src = qkv + (row0 + k) * 3 * d + h * DH;
    reinterpret_cast<__nv_bfloat162*>(K0 + k * slot_stride + (long long)pos * DH)[lane] =
        reinterpret_cast<const __nv_bfloat162*>(src + d)[lane];
    reinterpret_cast<__nv_bfloat162*>(V0 + k * slot_stride + (long long)pos * DH)[lane] =
        reinterpret_cast<const __nv_bfloat162*>(src + 2 * d)[lane];
    const float2 qq = __bfloat1622float2(reinterpret_cast<const __nv_bfloat162*>(src)[lane]);
    qs[k * DH + 2 * lane] = qq.x * 0.125f;
    qs[k * DH + 2 * lane + 1] = qq.y * 0.125f;
  }
  __syncwarp();
  const uint8_t* ind = indir + row0 * n_ctx;   // ind[k * n_ctx + j]
  // common prefix: position pos is always private (every row just appended to its own slot)
  int c = pos;
  for (int j = lane; j < pos; j += 32) {
    const int s0 = ind[j];
    bool same = true;
#pragma unroll
    for (int k = 1; k < K; ++k) same = same && ind[k * n_ctx + j] == s0;
    if (!same) { c = j; break; }
  }
  c = __reduce_min_sync(FULL_MASK, c);
  const int jg = lane >> 3, c8 = lane & 7;
  // ---- scores, shared prefix: one key row per 8 lanes, used by all K queries
  {
    float qf[K][8];
#pragma unroll
    for (int k = 0; k < K; ++k)
#pragma unroll
      for (int i = 0; i < 8; ++i) qf[k][i] = qs[k * DH + c8 * 8 + i];
#pragma unroll 2
    for (int j0 = 0; j0 < c; j0 += 4) {
      const int j = j0 + jg;
      const bool valid = j < c;
      uint4 u = make_uint4(0, 0, 0, 0);
      if (valid) u = reinterpret_cast<const uint4*>(K0 + (int)ind[j] * slot_stride)[(long long)j * 8 + c8];
      const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
      float f[8];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 t2 = __bfloat1622float2(hh[e]);
        f[2 * e] = t2.x;
        f[2 * e + 1] = t2.y;
      }
#pragma unroll
      for (int k = 0; k < K; ++k) {
        float sc = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) sc = fmaf(f[i], qf[k][i], sc);
        sc += __shfl_xor_sync(FULL_MASK, sc, 1);
        sc += __shfl_xor_sync(FULL_MASK, sc, 2);
        sc += __shfl_xor_sync(FULL_MASK, sc, 4);
        if (c8 == 0 && valid) ps[k * n_ctx + j] = sc;
      }
    }
    // ---- scores, private tail: per beam
#pragma unroll
    for (int k = 0; k < K; ++k) {
      for (int j0 = c; j0 <= pos; j0 += 4) {
        const int j = j0 + jg;
        const bool valid = j <= pos;
        uint4 u = make_uint4(0, 0, 0, 0);
        if (valid) u = reinterpret_cast<const uint4*>(K0 + (int)ind[k * n_ctx + j] * slot_stride)[(long long)j * 8 + c8];
        const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
        float sc = 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 t2 = __bfloat1622float2(hh[e]);
          sc = fmaf(t2.x, qf[k][2 * e], sc);
          sc = fmaf(t2.y, qf[k][2 * e + 1], sc);
        }
        sc += __shfl_xor_sync(FULL_MASK, sc, 1);
        sc += __shfl_xor_sync(FULL_MASK, sc, 2);
        sc += __shfl_xor_sync(FULL_MASK, sc, 4);
        if (c8 == 0 && valid) ps[k * n_ctx + j] = sc;
      }
    }
  }
  __syncwarp();
  // ---- softmax per beam
  float inv[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    float* pk = ps + k * n_ctx;
    float mx = -INFINITY;
    for (int j = lane; j <= pos; j += 32) mx = fmaxf(mx, pk[j]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j <= pos; j += 32) {
      const float p = __expf(pk[j] - mx);
      pk[j] = p;
      sum += p;
    }
    inv[k] = 1.f / warp_sum(sum);
  }
  __syncwarp();
  // ---- PV: shared prefix rows once for all beams, private tail per beam
  float o[K][8];
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int i = 0; i < 8; ++i) o[k][i] = 0.f;
#pragma unroll 2
  for (int j = jg; j < c; j += 4) {
    const uint4 u = reinterpret_cast<const uint4*>(V0 + (int)ind[j] * slot_stride)[(long long)j * 8 + c8];
    const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
    float f[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 t2 = __bfloat1622float2(hh[e]);
      f[2 * e] = t2.x;
      f[2 * e + 1] = t2.y;
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const float p = ps[k * n_ctx + j];
#pragma unroll
      for (int i = 0; i < 8; ++i) o[k][i] = fmaf(p, f[i], o[k][i]);
    }
  }
#pragma unroll
  for (int k = 0; k < K; ++k) {
    for (int j = c + jg; j <= pos; j += 4) {
      const uint4 u = reinterpret_cast<const uint4*>(V0 + (int)ind[k * n_ctx + j] * slot_stride)[(long long)j * 8 + c8];
      const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
      const float p = ps[k * n_ctx + j];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 t2 = __bfloat1622float2(hh[e]);
        o[k][2 * e] = fmaf(p, t2.x, o[k][2 * e]);
        o[k][2 * e + 1] = fmaf(p, t2.y, o[k][2 * e + 1]);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < K; ++k) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      o[k][i] += __shfl_xor_sync(FULL_MASK, o[k][i], 8);
      o[k][i] += __shfl_xor_sync(FULL_MASK, o[k][i], 16);
    }
    if (jg == 0) {
      uint4 wv;
      __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&wv);
#pragma unroll
      for (int e = 0; e < 4; ++e) hh[e] = __floats2bfloat162_rn(o[k][2 * e] * inv[k], o[k][2 * e + 1] * inv[k]);
      reinterpret_cast<uint4*>(out + (row0 + k) * d + h * DH)[c8] = wv;
    }
  }
}

template <int K>
static int launch_self_attn_beam(const __nv_bfloat16* qkv, __nv_bfloat16* kc, __nv_bfloat16* vc, const int32_t* pos, int d,
                                 int n_ctx, int H, int n_windows, __nv_bfloat16* out, const uint8_t* indir, cudaStream_t st) {
  const size_t smem = (size_t)SB_WARPS * K * (n_ctx + DH) * sizeof(float);
  static size_t attr[64] = {};
  const int dev = current_device_slot();
  if (smem > attr[dev] && smem > 48 * 1024) {
    JANUS_CUDA(cudaFuncSetAttribute(dec_self_attn_beam_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr[dev] = smem;
  }
  const int n_items = n_windows * H;
  JANUS_CUDA(launch_pdl(dec_self_attn_beam_kernel<K>, dim3((n_items + SB_WARPS - 1) / SB_WARPS), dim3(SB_WARPS * 32), smem, st,
                        qkv, kc, vc, pos, d, n_ctx, H, n_items, out, indir));
  return JANUS_OK;
}

// *handled = false: no kernel for this beam width (the caller falls back to dec_self_attn_kernel<true>)
static int dec_self_attn_beam(const __nv_bfloat16* qkv, __nv_bfloat16* kc, __nv_bfloat16* vc, const int32_t* pos, int d,
                              int n_ctx, int H, int n_windows, int beam, __nv_bfloat16* out, const uint8_t* indir,
                              cudaStream_t st, bool* handled) {
  *handled = true;
  switch (beam) {
    case 2: return launch_self_attn_beam<2>(qkv, kc, vc, pos, d, n_ctx, H, n_windows, out, indir, st);
    case 3: return launch_self_attn_beam<3>(qkv, kc, vc, pos, d, n_ctx, H, n_windows, out, indir, st);
    case 4: return launch_self_attn_beam<4>(qkv, kc, vc, pos, d, n_ctx, H, n_windows, out, indir, st);
    case 5: return launch_self_attn_beam<5>(qkv, kc, vc, pos, d, n_ctx, H, n_windows, out, indir, st);
    case 6: return launch_self_attn_beam<6>(qkv, kc, vc, pos, d, n_ctx, H, n_windows, out, indir, st);
    case 7: return launch_self_attn_beam<7>(qkv, kc, vc, pos, d, n_ctx, H, n_windows, out, indir, st);
    case 8: return launch_self_attn_beam<8>(qkv, kc, vc, pos, d, n_ctx, H, n_windows, out, indir, st);
    default: *handled = false; return JANUS_OK;
  }
}

// ------------------------------------------------------------------------------------------
// cross-attention over the encoder states: one CTA per (head, stream); HBM-bound
// ------------------------------------------------------------------------------------------
constexpr int XA_THREADS = 192;  // 192 * 8 keys = 1536 >= 1500

__global__ void __launch_bounds__(XA_THREADS)
dec_cross_attn_kernel(const __nv_bfloat16* q, const __nv_bfloat16* __restrict__ kT,
                      const __nv_bfloat16* __restrict__ v, int T, int kv_ld, int d, int H, int n_items,
                      __nv_bfloat16* out) {
  // 6.4 KB of shared memory and 40 registers per thread: with the grid capped at 3 CTAs per SM (xa_grid)
  // three of these fit beside one persistent tcgen05 GEMM CTA of the encoder stream (198 KB, 41 K registers)
  __shared__ float qs[DH];
  __shared__ __align__(16) float ps[XA_THREADS * 8];
  __shared__ float wred[2][XA_THREADS / 32];
  float (*red)[DH] = reinterpret_cast<float (*)[DH]>(ps);   // [XA_THREADS / 8][DH], reuses ps after the PV pass
  const int tid = threadIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
  const int b = item / H, h = item - b * H;
  const long long bh = (long long)b * H + h;
  if (tid < DH) qs[tid] = __bfloat162float(q[(long long)b * d + h * DH + tid]) * (0.125f * 1.4426950408889634f);
  __syncthreads();
  // ---- scores for keys 8*tid .. 8*tid+7
  const int key0 = tid * 8;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  if (key0 < T) {   // (threads whose 8 keys are all padding load nothing)
    const uint4* kp = reinterpret_cast<const uint4*>(kT + bh * DH * kv_ld + key0);
    const int stride = kv_ld >> 3;  // uint4 per dim row
#pragma unroll 16
    for (int dim = 0; dim < DH; ++dim) {
      const uint4 u = __ldg(kp + (long long)dim * stride);
      const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
      const float qd = qs[dim];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(hh[e]);
        acc[2 * e] = fmaf(f.x, qd, acc[2 * e]);
        acc[2 * e + 1] = fmaf(f.y, qd, acc[2 * e + 1]);
      }
    }
  }
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    acc[i] = (key0 + i < T) ? acc[i] : -INFINITY;  // select, not multiply: the pad keys are uninitialised
    mx = fmaxf(mx, acc[i]);
  }
  mx = warp_max(mx);
  if ((tid & 31) == 0) wred[0][tid >> 5] = mx;
  __syncthreads();
  mx = wred[0][0];
#pragma unroll
  for (int w = 1; w < XA_THREADS / 32; ++w) mx = fmaxf(mx, wred[0][w]);
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float p = exp2f(acc[i] - mx);
    ps[key0 + i] = p;
    sum += p;
  }
  sum = warp_sum(sum);
  if ((tid & 31) == 0) wred[1][tid >> 5] = sum;
  __syncthreads();
  sum = 0.f;
#pragma unroll
  for (int w = 0; w < XA_THREADS / 32; ++w) sum += wred[1][w];
  // ---- PV: 8 lanes per 128-byte V row, 24 rows per pass
  const int c = tid & 7, rg = tid >> 3;
  float o[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i] = 0.f;
  const uint4* vp = reinterpret_cast<const uint4*>(v + bh * (long long)T * DH) + c;
#pragma unroll 8
  for (int j = rg; j < T; j += XA_THREADS / 8) {
    const uint4 u = __ldg(vp + (long long)j * 8);
    const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
    const float p = ps[j];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(hh[e]);
      o[2 * e] = fmaf(p, f.x, o[2 * e]);
      o[2 * e + 1] = fmaf(p, f.y, o[2 * e + 1]);
    }
  }
  __syncthreads();   // every thread is done reading ps: reuse it for the cross-row reduction
#pragma unroll
  for (int i = 0; i < 8; ++i) red[rg][c * 8 + i] = o[i];
  __syncthreads();
  if (tid < DH) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < XA_THREADS / 8; ++r) t += red[r][tid];
    out[(long long)b * d + h * DH + tid] = __float2bfloat16_rn(t / sum);
  }
  }  // item loop (the next iteration's first __syncthreads orders the reuse of qs / ps)
}

// Grid of the cross-attention kernel: one CTA per (stream, head) by default; JANUS_XA_CTAS_PER_SM=n makes it
// persistent with n CTAs per SM so that tcgen05 CTAs of a concurrent encoder stream always find room.
static int xa_grid(int n_items) {
  const int cap = option(JANUS_OPT_XA_CTAS_PER_SM, 0);   // > 0: cap the cross-attention grid (overlap experiments)
  if (cap <= 0) return n_items;
  const long long g = (long long)cap * num_sms();
  return n_items < g ? n_items : (int)g;
}

// ------------------------------------------------------------------------------------------
// cross-attention for beam search: the NQ beams of a stream attend to the same K / V, so one CTA per
// (stream, head) streams them ONCE and serves all NQ queries (NQ x the FMAs per byte; still HBM-bound: 64 dims x
// NQ FMAs per 16-byte load). Same layouts and thread mapping as dec_cross_attn_kernel.
// ------------------------------------------------------------------------------------------
// FMAs are issued as packed fp32 pairs (fma.rn.f32x2: half the issue slots of 40 scalar FMAs per 16-byte load; the
// kernel is issue- / latency-bound, not FMA-pipe-bound); US / UV = unroll of the two streaming loops (the compiler
// software-pipelines the loads across the unrolled bodies). Measured at 512 streams x 12 heads x 5 beams (2.37 GB per
// launch, scripts/gpu_xb_bench.py): scalar FMAs 3.6-3.7 TB/s; packed 4.57 (US / UV = 8 / 4) - 4.72 TB/s (16 / 8) at 96
// registers, 3 CTAs per SM; gathering a batch of loads up front 3.2-4.3; forcing 4 / 5 CTAs per SM through
// __launch_bounds__ 4.0 / 2.3 (spills), 2 CTAs with 140 registers 4.1; a larger shared-memory carve-out does not help;
// p stored as (p, p) pairs (no MOV per packed multiplier, twice the shared memory) 4.61; the same body inlined
// from a __device__ function 3.64 (110 more register moves around the 64-bit pairs): the schedule is fragile,
// which is why the tensor-core kernel below (kv_layout 1) is the one beam sessions use.
template <int NQ, int US, int UV>
__global__ void __launch_bounds__(XA_THREADS)
dec_cross_attn_beam_kernel(const __nv_bfloat16* q, const __nv_bfloat16* __restrict__ kT,
                           const __nv_bfloat16* __restrict__ v, int T, int kv_ld, int d, int H, int n_items,
                           __nv_bfloat16* out) {
  extern __shared__ __align__(16) float xb_smem[];
  float (*ps)[XA_THREADS * 8] = reinterpret_cast<float (*)[XA_THREADS * 8]>(xb_smem);   // [NQ][1536]
  __shared__ __align__(8) float2 qs[NQ][DH];   // (q, q): one 8-byte load gives the packed multiplier
  __shared__ float wred[2][NQ][XA_THREADS / 32];
  const int tid = threadIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
  const int b = item / H, h = item - b * H;   // b: stream; its rows are b * NQ .. b * NQ + NQ - 1
  const long long bh = (long long)b * H + h;
  for (int i = tid; i < NQ * DH; i += XA_THREADS) {
    const int qi = i / DH, e = i - qi * DH;
    const float x = __bfloat162float(q[((long long)b * NQ + qi) * d + h * DH + e]) * (0.125f * 1.4426950408889634f);
    qs[qi][e] = make_float2(x, x);
  }
  __syncthreads();
  const int key0 = tid * 8;
  uint64_t acc2[NQ][4];
#pragma unroll
  for (int qi = 0; qi < NQ; ++qi)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc2[qi][i] = 0ull;
  if (key0 < T) {
    const uint4* kp = reinterpret_cast<const uint4*>(kT + bh * DH * kv_ld + key0);
    const int stride = kv_ld >> 3;
#pragma unroll US
    for (int dim = 0; dim < DH; ++dim) {
      const uint4 u = __ldg(kp + (long long)dim * stride);
      const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
      uint64_t f2[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 t2 = __bfloat1622float2(hh[e]);
        f2[e] = pack_f32x2(t2.x, t2.y);
      }
#pragma unroll
      for (int qi = 0; qi < NQ; ++qi) {
        const uint64_t qd = *reinterpret_cast<const uint64_t*>(&qs[qi][dim]);
#pragma unroll
        for (int e = 0; e < 4; ++e) acc2[qi][e] = fma_f32x2(f2[e], qd, acc2[qi][e]);
      }
    }
  }
#pragma unroll
  for (int qi = 0; qi < NQ; ++qi) {
    float a[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) unpack_f32x2(acc2[qi][e], a[2 * e], a[2 * e + 1]);
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a[i] = (key0 + i < T) ? a[i] : -INFINITY;
      mx = fmaxf(mx, a[i]);
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) acc2[qi][e] = pack_f32x2(a[2 * e], a[2 * e + 1]);
    mx = warp_max(mx);
    if ((tid & 31) == 0) wred[0][qi][tid >> 5] = mx;
  }
  __syncthreads();
#pragma unroll
  for (int qi = 0; qi < NQ; ++qi) {
    float m = wred[0][qi][0];
#pragma unroll
    for (int w = 1; w < XA_THREADS / 32; ++w) m = fmaxf(m, wred[0][qi][w]);
    float sm = 0.f;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float a0, a1;
      unpack_f32x2(acc2[qi][e], a0, a1);
      const float p0 = exp2f(a0 - m), p1 = exp2f(a1 - m);
      *reinterpret_cast<float2*>(&ps[qi][key0 + 2 * e]) = make_float2(p0, p1);
      sm += p0;
      sm += p1;
    }
    sm = warp_sum(sm);
    if ((tid & 31) == 0) wred[1][qi][tid >> 5] = sm;
  }
  __syncthreads();
  // ---- PV: 8 lanes per 128-byte V row, 24 rows per pass, every row used by all NQ queries
  const int c = tid & 7, rg = tid >> 3;
  uint64_t o2[NQ][4];
#pragma unroll
  for (int qi = 0; qi < NQ; ++qi)
#pragma unroll
    for (int i = 0; i < 4; ++i) o2[qi][i] = 0ull;
  const uint4* vp = reinterpret_cast<const uint4*>(v + bh * (long long)T * DH) + c;
#pragma unroll UV
  for (int j = rg; j < T; j += XA_THREADS / 8) {
    const uint4 u = __ldg(vp + (long long)j * 8);
    const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&u);
    uint64_t f2[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 t2 = __bfloat1622float2(hh[e]);
      f2[e] = pack_f32x2(t2.x, t2.y);
    }
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi) {
      const float p = ps[qi][j];
      const uint64_t pp = pack_f32x2(p, p);
#pragma unroll
      for (int e = 0; e < 4; ++e) o2[qi][e] = fma_f32x2(f2[e], pp, o2[qi][e]);
    }
  }
  __syncthreads();   // every thread is done reading ps: reuse each query's region for its cross-row reduction
#pragma unroll
  for (int qi = 0; qi < NQ; ++qi)
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float o0, o1;
      unpack_f32x2(o2[qi][e], o0, o1);
      *reinterpret_cast<float2*>(&ps[qi][rg * DH + c * 8 + 2 * e]) = make_float2(o0, o1);
    }
  __syncthreads();
  for (int i = tid; i < NQ * DH; i += XA_THREADS) {
    const int qi = i / DH, e = i - qi * DH;
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < XA_THREADS / 8; ++r) t += ps[qi][r * DH + e];
    float sm = 0.f;
#pragma unroll
    for (int w = 0; w < XA_THREADS / 32; ++w) sm += wred[1][qi][w];
    out[((long long)b * NQ + qi) * d + h * DH + e] = __float2bfloat16_rn(t / sm);
  }
  __syncthreads();   // qs / ps / wred are rewritten by the next item
  }
}

template <int NQ, int US, int UV>
static int launch_cross_attn_beam(const __nv_bfloat16* q, const __nv_bfloat16* kT, const __nv_bfloat16* v, int T, int kv_ld,
                                  int d, int H, int n_streams, __nv_bfloat16* out, cudaStream_t st) {
  const size_t smem = (size_t)NQ * XA_THREADS * 8 * sizeof(float);
  static bool attr[64] = {};
  const int dev = current_device_slot();
  if (!attr[dev]) {
    JANUS_CUDA(cudaFuncSetAttribute(dec_cross_attn_beam_kernel<NQ, US, UV>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)smem));
    attr[dev] = true;
  }
  JANUS_CUDA(launch_pdl(dec_cross_attn_beam_kernel<NQ, US, UV>, dim3(xa_grid(H * n_streams)), dim3(XA_THREADS), smem, st, q,
                        kT, v, T, kv_ld, d, H, H * n_streams, out));
  return JANUS_OK;
}

// ------------------------------------------------------------------------------------------
// cross-attention for beam search on the tensor cores (kv_layout 1): K row-major [n_ctx][64], V TRANSPOSED
// [64][kv_ld] (the two cache layouts of the greedy kernel with their roles swapped, written by the same GEMM
// epilogue). With these layouts both contractions are mma.sync.m16n8k16 whose B fragments are plain 16-byte
// global loads -- no shared-memory staging, no bf16 -> fp32 conversions, ~4x fewer instructions per byte than
// the FMA kernel above, which is what lets the kernel run at the HBM rate:
//   S[beams (rows 0-7 of the M = 16 tile; rows 8-15 are zero) x 8 keys] += Q[beams x 16 dims] K^T[16 dims x 8 keys]
//       lane (g = lane / 4, t = lane % 4) loads dims 16t .. 16t+15 of ITS key row (2 x 16 B): the contraction
//       index is permuted identically for Q and K, so consecutive dims feed consecutive k-steps
//   O[beams x 8 dims] += P[beams x 16 keys] V^T[16 keys x 8 dims]: lane (g, t) loads keys 8t .. 8t+7 of dim row g
//       (16 B); the S tiles of a 32-key block are assigned to key rows such that the lane already holds P for
//       exactly those 8 keys in its accumulator registers (column c of tile j <-> key 8 (c / 2) + 2 j + c % 2)
// One CTA per (stream, head), 4 warps; warp w takes the 32-key blocks w, w + 4, ... with a flash-style running
// maximum / sum, and the four partial results are merged through shared memory.
// ------------------------------------------------------------------------------------------
constexpr int XM_WARPS = 4;

__device__ __forceinline__ void xm_mma(float (&c)[4], uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
  // rows 8-15 of A (a1, a3) are zero: at most 8 beams
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(0u), "r"(a2), "r"(0u), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t xm_pack_bf16(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}

__global__ void __launch_bounds__(XM_WARPS * 32)
dec_cross_attn_beam_mma_kernel(const __nv_bfloat16* q, const __nv_bfloat16* __restrict__ kc,
                               const __nv_bfloat16* __restrict__ vt, int T, int kv_ld, int d, int H, int NQ,
                               __nv_bfloat16* out) {
  __shared__ float sm_m[XM_WARPS][8], sm_l[XM_WARPS][8];
  __shared__ float sm_o[XM_WARPS][8][DH];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  pdl_launch_dependents();
  pdl_wait();
  const int item = blockIdx.x;
  const int b = item / H, h = item - b * H;
  const long long bh = (long long)b * H + h;
  // Q fragments (unscaled bf16 straight from memory; the 1/8 log2(e) factor is applied to the fp32 scores)
  uint32_t qa0[4], qa2[4];
  {
    const uint32_t* qrow = reinterpret_cast<const uint32_t*>(q + ((long long)b * NQ + (g < NQ ? g : 0)) * d + h * DH + 16 * t);
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      qa0[s] = g < NQ ? qrow[2 * s] : 0u;       // dims 16t + 4s, +1
      qa2[s] = g < NQ ? qrow[2 * s + 1] : 0u;   // dims 16t + 4s + 2, +3
    }
  }
  const __nv_bfloat16* K = kc + bh * (long long)T * DH;
  const __nv_bfloat16* V = vt + bh * (long long)DH * kv_ld;
  const float sc = 0.125f * 1.4426950408889634f;
  float o[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt)
#pragma unroll
    for (int i = 0; i < 4; ++i) o[nt][i] = 0.f;
  float m_run = -INFINITY, l_run = 0.f;
  const int n_blk = (T + 31) >> 5;
#pragma unroll 1
  for (int blk = warp; blk < n_blk; blk += XM_WARPS) {
    const int base = blk << 5;
    // ---- loads: 8 x 16 B of K (4 key rows x 32 B) and 8 x 16 B of V^T (8 dim rows x 8 keys)
    uint4 ku[4][2], vu[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int row = base + 8 * (g >> 1) + 2 * j + (g & 1);
      row = row < T ? row : T - 1;   // clamped rows are masked below
      const uint4* kp = reinterpret_cast<const uint4*>(K + (long long)row * DH + 16 * t);
      ku[j][0] = __ldg(kp);
      ku[j][1] = __ldg(kp + 1);
    }
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
      vu[nt] = __ldg(reinterpret_cast<const uint4*>(V + (long long)(8 * nt + g) * kv_ld + base + 8 * t));
    // ---- scores of the 32 keys: tile j, column c <-> key base + 8 (c / 2) + 2 j + c % 2
    float c[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
      for (int i = 0; i < 4; ++i) c[j][i] = 0.f;
      xm_mma(c[j], qa0[0], qa2[0], ku[j][0].x, ku[j][0].y);
      xm_mma(c[j], qa0[1], qa2[1], ku[j][0].z, ku[j][0].w);
      xm_mma(c[j], qa0[2], qa2[2], ku[j][1].x, ku[j][1].y);
      xm_mma(c[j], qa0[3], qa2[3], ku[j][1].z, ku[j][1].w);
    }
    // this lane holds, for beam row g, the keys base + 8t + 2j + i (j = 0..3, i = 0..1) in c[j][i]
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int key = base + 8 * t + 2 * j + i;
        c[j][i] = key < T ? c[j][i] * sc : -INFINITY;
        mx = fmaxf(mx, c[j][i]);
      }
    mx = fmaxf(mx, __shfl_xor_sync(FULL_MASK, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(FULL_MASK, mx, 2));
    const float m_new = fmaxf(m_run, mx);   // finite: every block holds at least one valid key
    const float alpha = exp2f(m_run - m_new);
    m_run = m_new;
    l_run *= alpha;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) { o[nt][0] *= alpha; o[nt][1] *= alpha; }
    uint32_t pa[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float p0 = exp2f(c[j][0] - m_new), p1 = exp2f(c[j][1] - m_new);
      l_run += p0 + p1;
      pa[j] = xm_pack_bf16(p0, p1);
    }
    if (base + 32 > T) {   // tail block: V^T columns of keys >= T are not initialised (0 x NaN must not happen)
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        uint32_t* w = reinterpret_cast<uint32_t*>(&vu[nt]);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int key = base + 8 * t + 2 * e;
          if (key >= T) w[e] = 0u; else if (key + 1 >= T) w[e] &= 0x0000ffffu;
        }
      }
    }
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      xm_mma(o[nt], pa[0], pa[1], vu[nt].x, vu[nt].y);   // keys 8t + 0..3
      xm_mma(o[nt], pa[2], pa[3], vu[nt].z, vu[nt].w);   // keys 8t + 4..7
    }
  }
  l_run += __shfl_xor_sync(FULL_MASK, l_run, 1);
  l_run += __shfl_xor_sync(FULL_MASK, l_run, 2);
  if (t == 0) { sm_m[warp][g] = m_run; sm_l[warp][g] = l_run; }
#pragma unroll
  for (int nt = 0; nt < 8; ++nt)
    *reinterpret_cast<float2*>(&sm_o[warp][g][8 * nt + 2 * t]) = make_float2(o[nt][0], o[nt][1]);
  __syncthreads();
  for (int i = threadIdx.x; i < NQ * DH; i += XM_WARPS * 32) {
    const int r = i / DH, e = i - r * DH;
    float M = sm_m[0][r];
#pragma unroll
    for (int w = 1; w < XM_WARPS; ++w) M = fmaxf(M, sm_m[w][r]);
    float L = 0.f, acc = 0.f;
#pragma unroll
    for (int w = 0; w < XM_WARPS; ++w) {
      const float f = sm_m[w][r] == -INFINITY ? 0.f : exp2f(sm_m[w][r] - M);   // a warp without blocks (T < 128)
      L += sm_l[w][r] * f;
      acc += sm_o[w][r][e] * f;
    }
    out[((long long)b * NQ + r) * d + h * DH + e] = __float2bfloat16_rn(acc / L);
  }
}

// q / out: [n_streams * beam, d] (row = stream * beam + k); kT / v: one layer's cross K / V of n_streams streams
// (kv_layout 1: k is row-major [T][64] and v transposed [64][kv_ld] -> tensor-core kernel)
static int dec_cross_attn_beam(const __nv_bfloat16* q, const __nv_bfloat16* kT, const __nv_bfloat16* v, int T, int kv_ld,
                               int d, int H, int n_streams, int beam, __nv_bfloat16* out, cudaStream_t st,
                               int kv_layout = 0) {
  if (kv_layout == 1) {
    JANUS_CHECK_ARG(beam >= 1 && beam <= 8 && kv_ld % 32 == 0 && kv_ld >= ((T + 31) & ~31),
                    "cross_attn_beam (mma): beam=%d / kv_ld=%d unsupported", beam, kv_ld);
    JANUS_CUDA(launch_pdl(dec_cross_attn_beam_mma_kernel, dim3(H * n_streams), dim3(XM_WARPS * 32), 0, st, q, kT, v, T, kv_ld,
                          d, H, beam, out));
    return JANUS_OK;
  }
  switch (beam) {
    case 2: return launch_cross_attn_beam<2, 16, 8>(q, kT, v, T, kv_ld, d, H, n_streams, out, st);
    case 3: return launch_cross_attn_beam<3, 16, 8>(q, kT, v, T, kv_ld, d, H, n_streams, out, st);
    case 4: return launch_cross_attn_beam<4, 16, 8>(q, kT, v, T, kv_ld, d, H, n_streams, out, st);
    case 5: return launch_cross_attn_beam<5, 16, 8>(q, kT, v, T, kv_ld, d, H, n_streams, out, st);
    case 6: return launch_cross_attn_beam<6, 16, 8>(q, kT, v, T, kv_ld, d, H, n_streams, out, st);
    case 7: return launch_cross_attn_beam<7, 16, 8>(q, kT, v, T, kv_ld, d, H, n_streams, out, st);
    case 8: return launch_cross_attn_beam<8, 16, 8>(q, kT, v, T, kv_ld, d, H, n_streams, out, st);
    default: JANUS_CHECK_ARG(false, "cross_attn_beam: beam=%d unsupported", beam);
  }
  return JANUS_OK;
}

// ------------------------------------------------------------------------------------------
// logit rules + log-softmax + argmax: one CTA per stream, one pass over the fp32 logits
// ------------------------------------------------------------------------------------------
struct PickParams {
  const float* logits;  // [B, n_vocab_pad]
  const uint8_t* suppress;
  int32_t* tokens;
  const int32_t* forced;
  int32_t* choice;
  float* logprob;
  float* sum_logprob;
  int32_t* pos;
  uint8_t* done;
  int32_t* n_done;
  int n_vocab, n_vocab_pad, max_len, n_prompt, eot, blank, ts_begin, max_initial_ts;
};

struct Stat {  // running (max, argmax, sum exp(x - max))
  float m;
  int idx;
  float s;
};
__device__ __forceinline__ void stat_add(Stat& a, float x, int i) {
  if (x > a.m) {
    a.s = a.s * __expf(a.m - x) + 1.f;
    a.m = x;
    a.idx = i;
  } else {
    a.s += __expf(x - a.m);
  }
}
__device__ __forceinline__ void stat_merge(Stat& a, const Stat& o) {
  if (o.m == -INFINITY) return;
  if (a.m == -INFINITY) { a = o; return; }
  if (o.m > a.m || (o.m == a.m && o.idx < a.idx)) {
    a.s = a.s * __expf(a.m - o.m) + o.s;
    a.m = o.m;
    a.idx = o.idx;
  } else {
    a.s += o.s * __expf(o.m - a.m);
  }
}
__device__ __forceinline__ Stat stat_shfl(const Stat& a, int o) {
  Stat r;
  r.m = __shfl_xor_sync(FULL_MASK, a.m, o);
  r.idx = __shfl_xor_sync(FULL_MASK, a.idx, o);
  r.s = __shfl_xor_sync(FULL_MASK, a.s, o);
  return r;
}

constexpr int PICK_THREADS = 512;

__global__ void __launch_bounds__(PICK_THREADS) dec_pick_kernel(const PickParams p) {
  __shared__ Stat sh[2][PICK_THREADS / 32];
  const int b = blockIdx.x, tid = threadIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const int pos = *p.pos;        // index of the token just consumed; we choose token pos+1
  const int nxt = pos + 1;
  if (nxt >= p.max_len) return;
  int32_t* toks = p.tokens + (long long)b * p.max_len;
  if (nxt < p.n_prompt) return;  // still feeding the prompt
  const int n_gen = nxt - p.n_prompt;
  const int tb = p.ts_begin;
  // rule state from the already emitted tokens (ApplyTimestampRules)
  const int last = n_gen >= 1 ? toks[nxt - 1] : -1;
  const int penult = n_gen >= 2 ? toks[nxt - 2] : -1;
  const bool last_ts = n_gen >= 1 && last >= tb;
  const bool penult_ts = n_gen < 2 || penult >= tb;
  int ts_floor = tb;  // timestamps below this are forbidden
  {
    // the last timestamp emitted so far (they never decrease, so it is the max)
    int last_stamp = -1;
    for (int i = p.n_prompt; i < nxt; ++i) {
      const int t = toks[i];
      if (t >= tb) last_stamp = t;
    }
    if (last_stamp >= 0) ts_floor = (last_ts && !penult_ts) ? last_stamp : last_stamp + 1;
  }
  const float* lg = p.logits + (long long)b * p.n_vocab_pad;
  Stat text = {-INFINITY, 0x7fffffff, 0.f}, ts = {-INFINITY, 0x7fffffff, 0.f};
  // 4 tokens per thread per iteration: one 16-byte logit load + one 4-byte mask load
  const int first_lo = tb, first_hi = tb + p.max_initial_ts;
  for (int i0 = tid * 4; i0 < p.n_vocab; i0 += PICK_THREADS * 4) {
    const float4 x4 = *reinterpret_cast<const float4*>(lg + i0);   // rows are padded to a multiple of 64
    uint32_t sup4;
    if (i0 + 3 < p.n_vocab) {
      sup4 = *reinterpret_cast<const uint32_t*>(p.suppress + i0);
    } else {
      sup4 = 0x01010101u;
      for (int e = 0; e < 4 && i0 + e < p.n_vocab; ++e)
        sup4 = (sup4 & ~(0xffu << (8 * e))) | ((uint32_t)p.suppress[i0 + e] << (8 * e));
    }
    const float xs[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int i = i0 + e;
      bool ok = ((sup4 >> (8 * e)) & 0xffu) == 0;
      if (n_gen == 0) ok = ok && i != p.blank && i != p.eot && i >= first_lo && i <= first_hi;  // SuppressBlank + first timestamp <= 1.0 s
      if (last_ts) ok = ok && (penult_ts ? i < tb : i >= p.eot);          // timestamps come in pairs
      if (i >= tb) ok = ok && i >= ts_floor;                              // monotone timestamps
      if (!ok) continue;
      if (i < tb) stat_add(text, xs[e], i); else stat_add(ts, xs[e], i);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    stat_merge(text, stat_shfl(text, o));
    stat_merge(ts, stat_shfl(ts, o));
  }
  if ((tid & 31) == 0) { sh[0][tid >> 5] = text; sh[1][tid >> 5] = ts; }
  __syncthreads();
  if (tid == 0) {
    text = sh[0][0]; ts = sh[1][0];
    for (int w = 1; w < PICK_THREADS / 32; ++w) { stat_merge(text, sh[0][w]); stat_merge(ts, sh[1][w]); }
    const float lse_ts = ts.m == -INFINITY ? -INFINITY : ts.m + logf(ts.s);
    int tok;
    float lp;
    if (lse_ts > text.m) {  // timestamps carry more mass than any text token: sample a timestamp
      tok = ts.idx;
      lp = ts.m - lse_ts;
    } else {
      Stat all = text;
      stat_merge(all, ts);
      tok = all.idx;
      lp = -logf(all.s);
    }
    const bool was_done = p.done[b] != 0;
    if (p.choice) p.choice[(long long)b * p.max_len + nxt] = tok;
    int emit = tok;
    if (p.forced) {
      const int f = p.forced[(long long)b * p.max_len + nxt];
      if (f >= 0) emit = f;
    }
    if (was_done) { emit = p.eot; lp = 0.f; }
    toks[nxt] = emit;
    p.logprob[(long long)b * p.max_len + nxt] = lp;
    if (!was_done) {
      p.sum_logprob[b] += lp;
      if (emit == p.eot) { p.done[b] = 1; atomicAdd(p.n_done, 1); }
    }
  }
}

__global__ void dec_advance_kernel(int32_t* pos) {
  pdl_launch_dependents();
  pdl_wait();
  *pos += 1;
}

}  // namespace janus

using namespace janus;

// Skinny linear dispatch. Up to 64 streams the register-streaming mma.sync kernel has the shorter
// dependency chain for K <= 1024; for more streams (or long K / the vocabulary projection) the
// tcgen05 GEMM with 64-wide N tiles fills its 128-row MMA tiles and wins.
static int linear(const __nv_bfloat16* A, int K, const void* W, const float* bias, const __nv_bfloat16* residual,
                  void* out, int rows, int N, int act, int out_f32, cudaStream_t st, bool light = false) {
  if (rows <= 64 && (light || (K <= 1024 && N <= 8192)))
    return dec_linear(A, K, static_cast<const __nv_bfloat16*>(W), bias, residual, out, rows, N, K, act, out_f32, st,
                      /*pdl=*/true);
  GemmParams q = {};
  q.M_per_batch = rows; q.n_batch = 1; q.N = N; q.K = K;
  q.out = static_cast<__nv_bfloat16*>(out); q.ldc = N; q.out_batch_stride = (long long)rows * N;
  q.bias = bias; q.residual = residual; q.act = act; q.store_mode = out_f32 ? STORE_F32 : STORE_ROWMAJOR;
  // N tile: the widest of 256 / 128 / 64 that still yields about one CTA per SM (skinny M: many small tiles
  // beat few wide ones); at >= 512 rows the wide vocabulary / QKV / MLP projections fill the machine with 128-256
  const int m_tiles = (rows + 127) / 128;
  q.bn_max = 64;
  for (int bn = 256; bn > 64; bn >>= 1)
    if (N % bn == 0 && (long long)m_tiles * (N / bn) >= (num_sms() * 9) / 10) { q.bn_max = bn; break; }
  return gemm_bf16(A, K, (long long)rows * K, static_cast<const __nv_bfloat16*>(W), K, q, st);
}

extern "C" int32_t janus_dec_cross_attn_bf16(const void* q, const void* cross_k, const void* cross_v, void* out,
                                             int32_t batch, int32_t n_heads, int32_t n_ctx, int32_t kv_ld,
                                             void* stream) {
  JANUS_CHECK_ARG(q && cross_k && cross_v && out && batch > 0 && n_heads > 0, "cross_attn: bad argument");
  JANUS_CHECK_ARG(n_ctx <= XA_THREADS * 8 && kv_ld % 8 == 0 && kv_ld >= n_ctx && kv_ld <= XA_THREADS * 8,
                  "cross_attn: n_ctx=%d / kv_ld=%d unsupported", n_ctx, kv_ld);
  dec_cross_attn_kernel<<<xa_grid(n_heads * batch), XA_THREADS, 0, (cudaStream_t)stream>>>(
      static_cast<const __nv_bfloat16*>(q), static_cast<const __nv_bfloat16*>(cross_k),
      static_cast<const __nv_bfloat16*>(cross_v), n_ctx, kv_ld, n_heads * DH, n_heads, n_heads * batch,
      static_cast<__nv_bfloat16*>(out));
  JANUS_LAUNCH_CHECK("dec_cross_attn_kernel");
  return JANUS_OK;
}

extern "C" int32_t janus_dec_cross_attn_beam_bf16(const void* q, const void* cross_k, const void* cross_v, void* out,
                                                  int32_t n_streams, int32_t beam, int32_t n_heads, int32_t n_ctx,
                                                  int32_t kv_ld, int32_t kv_layout, void* stream) {
  JANUS_CHECK_ARG(kv_layout == 0 || kv_layout == 1, "cross_attn_beam: kv_layout must be 0 or 1");
  JANUS_CHECK_ARG(q && cross_k && cross_v && out && n_streams > 0 && n_heads > 0, "cross_attn_beam: bad argument");
  JANUS_CHECK_ARG(beam >= (kv_layout ? 1 : 2) && beam <= JANUS_MAX_BEAM, "cross_attn_beam: beam=%d must be in %d .. %d", beam,
                  kv_layout ? 1 : 2, JANUS_MAX_BEAM);
  JANUS_CHECK_ARG(n_ctx <= XA_THREADS * 8 && kv_ld % 8 == 0 && kv_ld >= n_ctx && kv_ld <= XA_THREADS * 8,
                  "cross_attn_beam: n_ctx=%d / kv_ld=%d unsupported", n_ctx, kv_ld);
  return dec_cross_attn_beam(static_cast<const __nv_bfloat16*>(q), static_cast<const __nv_bfloat16*>(cross_k),
                             static_cast<const __nv_bfloat16*>(cross_v), n_ctx, kv_ld, n_heads * DH, n_heads, n_streams,
                             beam, static_cast<__nv_bfloat16*>(out), (cudaStream_t)stream, kv_layout);
}

static int32_t cross_kv_impl(const janus_whisper_dims_t* c, const janus_decoder_weights_t* w, const void* enc_out,
                             int32_t batch, void* cross_k, void* cross_v, int kv_swap, void* stream) {
  JANUS_CHECK_ARG(c && w && enc_out && cross_k && cross_v && batch > 0, "cross_kv: bad argument");
  const int T = c->n_ctx, d = c->d;
  GemmParams p = {};
  p.M_per_batch = T; p.n_batch = batch; p.N = c->n_layers * 2 * d; p.K = d;
  p.bias = w->b_ckv; p.store_mode = STORE_CROSS_KV;
  p.kv_k = static_cast<__nv_bfloat16*>(cross_k); p.kv_v = static_cast<__nv_bfloat16*>(cross_v);
  p.kv_d = d; p.kv_heads = c->n_heads; p.kv_ld = w->kv_ld; p.kv_swap = kv_swap;
  p.ldc = 0; p.out_batch_stride = 0;
  return gemm_bf16(static_cast<const __nv_bfloat16*>(enc_out), d, (long long)T * d,
                   static_cast<const __nv_bfloat16*>(w->w_ckv), d, p, (cudaStream_t)stream);
}

extern "C" int32_t janus_whisper_cross_kv_bf16(const janus_whisper_dims_t* c, const janus_decoder_weights_t* w,
                                               const void* enc_out, int32_t batch, void* cross_k, void* cross_v,
                                               void* stream) {
  return cross_kv_impl(c, w, enc_out, batch, cross_k, cross_v, 0, stream);
}

extern "C" int32_t janus_whisper_cross_kv_swapped_bf16(const janus_whisper_dims_t* c, const janus_decoder_weights_t* w,
                                                       const void* enc_out, int32_t batch, void* cross_k, void* cross_v,
                                                       void* stream) {
  return cross_kv_impl(c, w, enc_out, batch, cross_k, cross_v, 1, stream);
}

// One decode step for s->batch rows. bm == nullptr: greedy (a row is a stream). bm != nullptr: beam search, rows are
// the beams of bm->n_streams streams (row = stream * beam + k): self-attention follows the beam indirection, the
// cross-attention of a stream's beams reads the same encoder states / K/V, and the token choice is the candidate /
// update pair of dec_beam.cu.
static int32_t decode_step_impl(const janus_whisper_dims_t* c, const janus_decoder_weights_t* w,
                                const janus_decode_state_t* s, const janus_beam_state_t* bm, void* stream_) {
  JANUS_CHECK_ARG(c && w && s, "decode_step: null pointer");
  const int grp = bm ? bm->beam : 1;
  JANUS_CHECK_ARG(!bm || !(s->flags & JANUS_DECODE_LIGHT), "decode_beam_step: LIGHT sessions are greedy only");
  JANUS_CHECK_ARG(c->d == c->n_heads * 64, "decode_step: head_dim must be 64");
  JANUS_CHECK_ARG(c->n_ctx <= XA_THREADS * 8 && w->kv_ld % 8 == 0 && w->kv_ld >= c->n_ctx && w->kv_ld <= XA_THREADS * 8,
                  "decode_step: n_ctx=%d / kv_ld=%d unsupported", c->n_ctx, w->kv_ld);
  JANUS_CHECK_ARG(w->n_text_ctx <= 512, "decode_step: n_text_ctx must be <= 512");
  const bool light = (s->flags & JANUS_DECODE_LIGHT) != 0;
  const bool latent = (s->flags & JANUS_DECODE_LATENT) != 0;
  JANUS_CHECK_ARG(!latent || (!light && s->enc_out && s->qlat && s->zlat),
                  "decode_step: JANUS_DECODE_LATENT needs enc_out / qlat / zlat and excludes LIGHT");
  JANUS_CHECK_ARG(latent || (s->cross_k && s->cross_v), "decode_step: cross_k / cross_v missing");
  JANUS_CHECK_ARG(!light || s->batch <= 64, "decode_step: JANUS_DECODE_LIGHT needs batch <= 64 (got %d)", s->batch);
  cudaStream_t st = (cudaStream_t)stream_;
  const int B = s->batch, d = c->d, H = c->n_heads, T = c->n_ctx;
  __nv_bfloat16* x = static_cast<__nv_bfloat16*>(s->x);
  __nv_bfloat16* xn = static_cast<__nv_bfloat16*>(s->xn);
  __nv_bfloat16* qkv = static_cast<__nv_bfloat16*>(s->qkv);
  __nv_bfloat16* att = static_cast<__nv_bfloat16*>(s->att);
  __nv_bfloat16* q = static_cast<__nv_bfloat16*>(s->q);
  __nv_bfloat16* ff = static_cast<__nv_bfloat16*>(s->ff);
  int rc;
  JANUS_CUDA(launch_pdl(dec_embed_kernel, dim3(B), dim3(128), 0, st, s->tokens, s->max_len, s->pos,
                        static_cast<const __nv_bfloat16*>(w->tok_emb), w->pos, d, x));
  const int self_ld = s->self_ld > 0 ? s->self_ld : w->n_text_ctx;
  JANUS_CHECK_ARG(self_ld >= s->max_len && self_ld <= 512, "decode_step: self_ld=%d must be in [max_len=%d, 512]", self_ld,
                  s->max_len);
  const long long self_stride = (long long)B * H * self_ld * DH;
  const bool kv_swapped = (s->flags & JANUS_DECODE_KV_SWAPPED) != 0;
  JANUS_CHECK_ARG(!kv_swapped || (bm && !latent), "decode_step: JANUS_DECODE_KV_SWAPPED is a beam-search K/V layout");
  const long long kt_stride = (long long)(B / grp) * H * DH * w->kv_ld;   // cross K/V exist once per stream
  const long long rm_stride = (long long)(B / grp) * H * T * DH;
  const long long ck_stride = kv_swapped ? rm_stride : kt_stride;
  const long long cv_stride = kv_swapped ? kt_stride : rm_stride;
  for (int l = 0; l < c->n_layers; ++l) {
    const janus_dec_layer_t& L = w->layers[l];
    if ((rc = layernorm_bf16(x, L.ln1_w, L.ln1_b, xn, B, d, d, d, st, /*pdl=*/true))) return rc;
    if ((rc = linear(xn, d, L.w_qkv, L.b_qkv, nullptr, qkv, B, 3 * d, ACT_NONE, 0, st, light))) return rc;
    // JANUS_OPT_BEAM_SHARED_SELF_ATTN = 0: per-row kernel through the indirection instead of the shared-prefix one
    const int beam_sa = option(JANUS_OPT_BEAM_SHARED_SELF_ATTN, 1);
    bool sa_done = false;
    if (bm && beam_sa &&
        (rc = dec_self_attn_beam(qkv, static_cast<__nv_bfloat16*>(s->self_k) + l * self_stride,
                                 static_cast<__nv_bfloat16*>(s->self_v) + l * self_stride, s->pos, d, self_ld, H, B / grp, grp,
                                 att, bm->indir, st, &sa_done)))
      return rc;
    if (sa_done) {
    } else if (bm) {
      JANUS_CUDA(launch_pdl(dec_self_attn_kernel<true>, dim3((H * B + SA_WARPS - 1) / SA_WARPS), dim3(SA_WARPS * 32), 0, st,
                            qkv, static_cast<__nv_bfloat16*>(s->self_k) + l * self_stride,
                            static_cast<__nv_bfloat16*>(s->self_v) + l * self_stride, s->pos, d, self_ld, H, H * B, att,
                            bm->indir, bm->beam));
    } else {
    JANUS_CUDA(launch_pdl(dec_self_attn_kernel<false>, dim3((H * B + SA_WARPS - 1) / SA_WARPS), dim3(SA_WARPS * 32), 0, st, qkv,
                          static_cast<__nv_bfloat16*>(s->self_k) + l * self_stride,
                          static_cast<__nv_bfloat16*>(s->self_v) + l * self_stride, s->pos, d, self_ld, H, H * B, att,
                          static_cast<const uint8_t*>(nullptr), 1));
    }
    if ((rc = linear(att, d, L.w_o, L.b_o, x, x, B, d, ACT_NONE, 0, st, light))) return rc;
    if ((rc = layernorm_bf16(x, L.lnc_w, L.lnc_b, xn, B, d, d, d, st, /*pdl=*/true))) return rc;
    if ((rc = linear(xn, d, L.w_cq, L.b_cq, nullptr, q, B, d, ACT_NONE, 0, st, light))) return rc;
    if (latent) {
      JANUS_CHECK_ARG(L.w_ck_lat && L.w_cv && L.b_cv, "decode_step: layer %d has no latent cross-attention weights", l);
      __nv_bfloat16* qlat = static_cast<__nv_bfloat16*>(s->qlat);
      __nv_bfloat16* zlat = static_cast<__nv_bfloat16*>(s->zlat);
      {   // absorbed queries: qlat[b, h, :] = q[b, h*64 .. h*64+63] . Wk_h   (grouped over heads, K = 64)
        GemmParams g = {};
        g.M_per_batch = B; g.n_batch = H; g.N = d; g.K = DH;
        g.out = qlat; g.ldc = 16LL * d; g.out_batch_stride = d;
        g.store_mode = STORE_ROWMAJOR; g.w_batch_rows = d;
        if ((rc = gemm_bf16(q, d, DH, static_cast<const __nv_bfloat16*>(L.w_ck_lat), DH, g, st))) return rc;
      }
      if ((rc = dec_xlat(qlat, static_cast<const __nv_bfloat16*>(s->enc_out), zlat, B, T, d, H, st, grp))) return rc;
      {   // att[b, h*64 .. h*64+63] = zlat[b, h, :] . Wv_h^T + bv_h   (grouped over heads, N = 64)
        GemmParams g = {};
        g.M_per_batch = B; g.n_batch = H; g.N = DH; g.K = d;
        g.out = att; g.ldc = d; g.out_batch_stride = DH;
        g.store_mode = STORE_ROWMAJOR; g.w_batch_rows = DH;
        g.bias = L.b_cv; g.bias_batch_stride = DH;
        if ((rc = gemm_bf16(zlat, (long long)H * d, d, static_cast<const __nv_bfloat16*>(L.w_cv), d, g, st))) return rc;
      }
    } else {
    if (grp > 1 || kv_swapped) {   // beam search: one CTA per (stream, head) serves all beams from one pass over K / V
      if ((rc = dec_cross_attn_beam(q, static_cast<const __nv_bfloat16*>(s->cross_k) + l * ck_stride,
                                    static_cast<const __nv_bfloat16*>(s->cross_v) + l * cv_stride, T, w->kv_ld, d, H,
                                    B / grp, grp, att, st, kv_swapped ? 1 : 0))) return rc;
    } else {
    JANUS_CUDA(launch_pdl(dec_cross_attn_kernel, dim3(xa_grid(H * B)), dim3(XA_THREADS), 0, st, q,
                          static_cast<const __nv_bfloat16*>(s->cross_k) + l * ck_stride,
                          static_cast<const __nv_bfloat16*>(s->cross_v) + l * cv_stride, T, w->kv_ld, d, H, H * B, att));
    }
    }
    if ((rc = linear(att, d, L.w_co, L.b_co, x, x, B, d, ACT_NONE, 0, st, light))) return rc;
    if ((rc = layernorm_bf16(x, L.ln2_w, L.ln2_b, xn, B, d, d, d, st, /*pdl=*/true))) return rc;
    if ((rc = linear(xn, d, L.w_fc1, L.b_fc1, nullptr, ff, B, 4 * d, ACT_GELU, 0, st, light))) return rc;
    if ((rc = linear(ff, 4 * d, L.w_fc2, L.b_fc2, x, x, B, d, ACT_NONE, 0, st, light))) return rc;
  }
  if ((rc = layernorm_bf16(x, w->ln_w, w->ln_b, xn, B, d, d, d, st, /*pdl=*/true))) return rc;
  if ((rc = linear(xn, d, w->tok_emb, nullptr, nullptr, s->logits, B, w->n_vocab_pad, ACT_NONE, 1, st, light))) return rc;
  if (bm) {
    BeamStepParams bp = {};
    bp.logits = s->logits; bp.suppress = w->suppress_mask; bp.tokens = s->tokens; bp.pos = s->pos; bp.n_done = s->n_done;
    bp.n_vocab = w->n_vocab; bp.n_vocab_pad = w->n_vocab_pad; bp.max_len = s->max_len; bp.n_prompt = s->n_prompt;
    bp.eot = w->eot; bp.blank = w->blank; bp.ts_begin = w->timestamp_begin; bp.max_initial_ts = w->max_initial_timestamp_index;
    bp.self_ld = self_ld; bp.bm = *bm;
    if ((rc = dec_beam_step(bp, st))) return rc;
    JANUS_CUDA(launch_pdl(dec_advance_kernel, dim3(1), dim3(1), 0, st, s->pos));
    return JANUS_OK;
  }
  PickParams pp;
  pp.logits = s->logits; pp.suppress = w->suppress_mask; pp.tokens = s->tokens; pp.forced = s->forced;
  pp.choice = s->choice; pp.logprob = s->logprob; pp.sum_logprob = s->sum_logprob; pp.pos = s->pos;
  pp.done = s->done; pp.n_done = s->n_done; pp.n_vocab = w->n_vocab; pp.n_vocab_pad = w->n_vocab_pad;
  pp.max_len = s->max_len; pp.n_prompt = s->n_prompt; pp.eot = w->eot; pp.blank = w->blank;
  pp.ts_begin = w->timestamp_begin; pp.max_initial_ts = w->max_initial_timestamp_index;
  JANUS_CUDA(launch_pdl(dec_pick_kernel, dim3(B), dim3(PICK_THREADS), 0, st, pp));
  JANUS_CUDA(launch_pdl(dec_advance_kernel, dim3(1), dim3(1), 0, st, s->pos));
  return JANUS_OK;
}

extern "C" int32_t janus_whisper_decode_step(const janus_whisper_dims_t* c, const janus_decoder_weights_t* w,
                                             const janus_decode_state_t* s, void* stream) {
  return decode_step_impl(c, w, s, nullptr, stream);
}

extern "C" int32_t janus_whisper_decode_beam_step(const janus_whisper_dims_t* c, const janus_decoder_weights_t* w,
                                                  const janus_decode_state_t* s, const janus_beam_state_t* bm,
                                                  void* stream) {
  JANUS_CHECK_ARG(bm && s, "decode_beam_step: null pointer");
  JANUS_CHECK_ARG(bm->beam >= 1 && bm->beam <= JANUS_MAX_BEAM && bm->n_streams > 0 && s->batch == bm->n_streams * bm->beam,
                  "decode_beam_step: batch=%d must be n_streams=%d x beam=%d (beam <= %d)", s->batch, bm->n_streams,
                  bm->beam, JANUS_MAX_BEAM);
  JANUS_CHECK_ARG(bm->max_candidates >= 1 && bm->max_candidates <= 2 * JANUS_MAX_BEAM, "decode_beam_step: max_candidates=%d",
                  bm->max_candidates);
  JANUS_CHECK_ARG(bm->cum && bm->indir && bm->cand_score && bm->cand_tok && bm->fin_tokens && bm->fin_score && bm->fin_len &&
                  bm->n_fin && bm->stream_done, "decode_beam_step: beam state has null buffers");
  return decode_step_impl(c, w, s, bm, stream);
}

extern "C" int32_t janus_whisper_beam_finalize(const janus_decode_state_t* s, const janus_beam_state_t* bm, int32_t eot,
                                               int32_t* best_tokens, float* best_score, int32_t* best_len, void* stream) {
  JANUS_CHECK_ARG(s && bm && best_tokens && best_score && best_len, "beam_finalize: null pointer");
  return dec_beam_finalize(s->tokens, s->pos, s->max_len, s->n_prompt, eot, *bm, best_tokens, best_score, best_len,
                           (cudaStream_t)stream);
}
