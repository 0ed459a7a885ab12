// K4: YIN pitch + per-hop level + stream RMS for the Janus prosody path (sm_100a).
//
// Replaces aubio.pitch('yin', 4096, hop, sr) as driven by
// /root/reference/backend/services/prosody.py:32-34,76-87 (upstream aubio 0.4.9
// src/pitch/pitch.c, pitchyin.c, mathutils.c).
//
// Parity strategy: BIT-EXACT float32. aubio evaluates, per hop,
//     d(tau) = sum_{j=0}^{W-1} (x[j] - x[j+tau])^2        (float32, j ascending, no FMA)
//     run   += d(tau);  c(tau) = d(tau) * (tau / run)      (strictly sequential prefix sum)
//     exit at the first tau > 4 with c(tau-3) < tol and c(tau-3) < c(tau-2)
// Every d(tau) is an independent sequential sum, so we give each lane R=8 consecutive lags
// and let it walk j in ascending order with separately rounded sub / mul / add
// (__fsub_rn/__fmul_rn/__fadd_rn are never contracted into FMA). The 8 lags of a lane share a
// sliding register window of x[j+tau..], so shared memory traffic is 2 LDS.128 per 96 FP32 ops.
// The prefix sum is carried lane to lane in order, so run/c are bit-identical as well, and
// therefore so are the early-exit decision, the interpolated period and f0.
//
// Work decomposition: one warp = one frame (hop); it evaluates lags in blocks of 256 and stops at
// the block in which aubio's early exit fires (tol = 0.8 makes that the first block for almost
// every voiced frame). Hops that fail aubio's -50 dB silence gate skip YIN entirely (their f0 is
// forced to 0 by aubio_pitch_do regardless). One CTA = 8 consecutive frames of one stream, which
// share a (7*hop + buf)-sample window staged once in shared memory.
//
// The kernel is FP32-pipe bound by construction (3 dependent-rounding FP32 ops per (j, tau) pair);
// HBM traffic is ~1.9x the 4 B/sample algorithmic minimum and irrelevant to its run time.
#include "common.cuh"

namespace janus {

constexpr int YIN_F = 8;            // frames (= warps) per CTA
constexpr int YIN_R = 8;            // lags per lane
constexpr int YIN_TB = 32 * YIN_R;  // lags per warp pass

// 16-byte groups are XOR-swizzled so that lanes reading float4 at a stride of 8 floats
// (group stride 2) hit 8 distinct bank groups per quarter warp.
__device__ __forceinline__ int swz(int g) { return g ^ ((g >> 3) & 1); }
__device__ __forceinline__ int swz_word(int i) { return (swz(i >> 2) << 2) | (i & 3); }

struct YinParams {
  const float* pcm;
  const int64_t* stream_offsets;
  const int64_t* stream_lengths;
  const int64_t* hop_offsets;
  const int32_t* group_offsets;
  const float* state_in;
  float* f0_out;
  float* hop_db_out;
  float* hop_sumsq_out;
  int32_t n_streams;
  int32_t sample_rate;
  int32_t buf;
  int32_t hop;
  int32_t ns_alloc;  // floats of sample staging (padded)
  float tol;
  float silence_db;
};

#define YIN_ACC(k, xv, wv)                            \
  {                                                   \
    float t__ = __fsub_rn((xv), (wv));                \
    acc[k] = __fadd_rn(acc[k], __fmul_rn(t__, t__));  \
  }

// one j-chunk of 4 samples against the 8 lags of this lane. A,B,C = x[j+tau0 .. j+tau0+11].
#define YIN_STEP(A, B, C, D, s)                                                         \
  {                                                                                     \
    D = S4[swz(gw + (s) + 3)];                                                          \
    xn = S4[swz(gb0 + (s) + 1)];                                                        \
    YIN_ACC(0, xb.x, A.x) YIN_ACC(1, xb.x, A.y) YIN_ACC(2, xb.x, A.z) YIN_ACC(3, xb.x, A.w) \
    YIN_ACC(4, xb.x, B.x) YIN_ACC(5, xb.x, B.y) YIN_ACC(6, xb.x, B.z) YIN_ACC(7, xb.x, B.w) \
    YIN_ACC(0, xb.y, A.y) YIN_ACC(1, xb.y, A.z) YIN_ACC(2, xb.y, A.w) YIN_ACC(3, xb.y, B.x) \
    YIN_ACC(4, xb.y, B.y) YIN_ACC(5, xb.y, B.z) YIN_ACC(6, xb.y, B.w) YIN_ACC(7, xb.y, C.x) \
    YIN_ACC(0, xb.z, A.z) YIN_ACC(1, xb.z, A.w) YIN_ACC(2, xb.z, B.x) YIN_ACC(3, xb.z, B.y) \
    YIN_ACC(4, xb.z, B.z) YIN_ACC(5, xb.z, B.w) YIN_ACC(6, xb.z, C.x) YIN_ACC(7, xb.z, C.y) \
    YIN_ACC(0, xb.w, A.w) YIN_ACC(1, xb.w, B.x) YIN_ACC(2, xb.w, B.y) YIN_ACC(3, xb.w, B.z) \
    YIN_ACC(4, xb.w, B.w) YIN_ACC(5, xb.w, C.x) YIN_ACC(6, xb.w, C.y) YIN_ACC(7, xb.w, C.z) \
    xb = xn;                                                                            \
  }

// mathutils.c fvec_quadratic_peak_pos (float32, same operation order)
__device__ __forceinline__ float quadratic_peak_pos(const float* c, int W, int pos) {
  if (pos == 0 || pos == W - 1) return (float)pos;
  const float s0 = c[pos - 1], s1 = c[pos], s2 = c[pos + 1];
  const float num = __fmul_rn(0.5f, __fsub_rn(s0, s2));
  const float den = __fadd_rn(__fsub_rn(s0, __fmul_rn(2.f, s1)), s2);
  return __fadd_rn((float)pos, __fdiv_rn(num, den));
}

__global__ void __launch_bounds__(YIN_F * 32, 2) yin_frames_kernel(const YinParams p) {
  extern __shared__ __align__(16) float smem[];
  float* S = smem;
  const float4* S4 = reinterpret_cast<const float4*>(smem);
  const int W = p.buf >> 1;
  float* CW = smem + p.ns_alloc + (threadIdx.x >> 5) * W;

  // ---- locate (stream, first hop) of this group
  const int g = blockIdx.x;
  int lo = 0, hi = p.n_streams;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (p.group_offsets[mid] <= g) lo = mid; else hi = mid;
  }
  const int s = lo;
  const int64_t off = p.stream_offsets[s];
  const int64_t len = p.stream_lengths[s];
  const int64_t hop_base = p.hop_offsets[s];
  const int64_t n_hops = p.hop_offsets[s + 1] - hop_base;
  const int64_t h0 = (int64_t)(g - p.group_offsets[s]) * YIN_F;

  // ---- stage the virtual signal V = state ++ pcm ++ zeros, V[(h0+1)*hop .. +NS)
  const int NS = (YIN_F - 1) * p.hop + p.buf;
  const int64_t V0 = (h0 + 1) * (int64_t)p.hop;
  const float* st = p.state_in ? p.state_in + (int64_t)s * p.buf : nullptr;
  const float* src = p.pcm + off;
  for (int i = threadIdx.x; i < p.ns_alloc; i += blockDim.x) {
    float v = 0.f;
    if (i < NS) {
      const int64_t V = V0 + i;
      if (V < p.buf) {
        if (st) v = st[V];
      } else if (V - p.buf < len) {
        v = __ldg(src + (V - p.buf));
      }
    }
    S[swz_word(i)] = v;
  }
  __syncthreads();

  const int f = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int64_t h = h0 + f;
  if (h >= n_hops) return;

  // ---- aubio_level_lin on the incoming hop: sequential float32 (gate), plus a tree sum (RMS)
  const int hop_word0 = f * p.hop + p.buf - p.hop;
  float e = 0.f;
  for (int i = 0; i < p.hop; i += 4) {
    const float4 v = S4[swz((hop_word0 + i) >> 2)];
    e = __fadd_rn(e, __fmul_rn(v.x, v.x));
    e = __fadd_rn(e, __fmul_rn(v.y, v.y));
    e = __fadd_rn(e, __fmul_rn(v.z, v.z));
    e = __fadd_rn(e, __fmul_rn(v.w, v.w));
  }
  float ts = 0.f;
  for (int gi = lane; gi < (p.hop >> 2); gi += 32) {
    const float4 v = S4[swz((hop_word0 >> 2) + gi)];
    ts += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
  }
  ts = warp_sum(ts);
  const float db = 10.f * log10f(__fdiv_rn(e, (float)p.hop));
  const bool gated = db < p.silence_db;
  if (lane == 0) {
    p.hop_db_out[hop_base + h] = db;
    p.hop_sumsq_out[hop_base + h] = ts;
  }
  if (gated) {  // aubio_pitch_do zeroes the pitch whatever the detector said
    if (lane == 0) p.f0_out[hop_base + h] = 0.f;
    return;
  }

  // ---- YIN, 256 lags per pass
  const int gb0 = (f * p.hop) >> 2;
  const int nblk = W / YIN_TB;
  const int steps = W >> 2;
  float run = 0.f;  // warp-uniform running sum of d
  int pos = -1;
  for (int b = 0; b < nblk; ++b) {
    const int tau0 = b * YIN_TB + lane * YIN_R;
    const int gw = gb0 + (tau0 >> 2);
    float acc[YIN_R];
#pragma unroll
    for (int k = 0; k < YIN_R; ++k) acc[k] = 0.f;
    float4 q0 = S4[swz(gw)], q1 = S4[swz(gw + 1)], q2 = S4[swz(gw + 2)], q3;
    float4 xb = S4[swz(gb0)], xn;
#pragma unroll 1
    for (int stp = 0; stp < steps; stp += 4) {
      YIN_STEP(q0, q1, q2, q3, stp + 0)
      YIN_STEP(q1, q2, q3, q0, stp + 1)
      YIN_STEP(q2, q3, q0, q1, stp + 2)
      YIN_STEP(q3, q0, q1, q2, stp + 3)
    }
    // strictly ordered prefix sum: lane 0's 8 lags, then lane 1's, ...
    float pre[YIN_R];
#pragma unroll 1
    for (int l = 0; l < 32; ++l) {
      if (lane == l) {
#pragma unroll
        for (int k = 0; k < YIN_R; ++k) {
          run = __fadd_rn(run, acc[k]);
          pre[k] = run;
        }
      }
      run = __shfl_sync(FULL_MASK, run, l);
    }
#pragma unroll
    for (int k = 0; k < YIN_R; ++k) {
      const int tau = tau0 + k;
      float c = (pre[k] != 0.f) ? __fmul_rn(acc[k], __fdiv_rn((float)tau, pre[k])) : 1.f;
      if (tau == 0) c = 1.f;
      CW[tau] = c;
    }
    __syncwarp();
    // early exit: first p in this pass with c[p] < tol && c[p] < c[p+1]  (p = tau - 3, tau > 4)
    const int p_lo = b * YIN_TB - 1 + lane * YIN_R;
    int first = -1;
#pragma unroll
    for (int k = YIN_R - 1; k >= 0; --k) {
      const int pp = p_lo + k;
      if (pp >= 2 && pp <= W - 4) {
        const float c0 = CW[pp], c1 = CW[pp + 1];
        if (c0 < p.tol && c0 < c1) first = pp;
      }
    }
    const unsigned m = __ballot_sync(FULL_MASK, first >= 0);
    if (m) {
      pos = __shfl_sync(FULL_MASK, first, __ffs(m) - 1);
      break;
    }
  }
  if (pos < 0) {
    // fvec_min_elem: smallest value, LAST index on ties
    float best = CW[lane];
    int bi = lane;
    for (int i = lane + 32; i < W; i += 32) {
      const float v = CW[i];
      if (!(best < v)) { best = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(FULL_MASK, best, o);
      const int oi = __shfl_xor_sync(FULL_MASK, bi, o);
      if (ov < best || (ov == best && oi > bi)) { best = ov; bi = oi; }
    }
    pos = bi;
  }
  if (lane == 0) {
    const float period = quadratic_peak_pos(CW, W, pos);
    float f0 = 0.f;
    if (period > 0.f) f0 = (float)((double)p.sample_rate / ((double)period + 0.));
    p.f0_out[hop_base + h] = f0;
  }
}

__global__ void yin_state_kernel(const float* __restrict__ pcm, const int64_t* __restrict__ stream_offsets,
                                 const int64_t* __restrict__ stream_lengths, int buf, int hop,
                                 const float* __restrict__ state_in, float* __restrict__ state_out) {
  const int s = blockIdx.x;
  const int64_t len = stream_lengths[s];
  const int64_t n_hops = (len + hop - 1) / hop;
  const int64_t V0 = n_hops * hop;  // new state = V[V0 .. V0+buf)
  const float* st = state_in ? state_in + (int64_t)s * buf : nullptr;
  const float* src = pcm + stream_offsets[s];
  for (int i = threadIdx.x; i < buf; i += blockDim.x) {
    const int64_t V = V0 + i;
    float v = 0.f;
    if (V < buf) {
      if (st) v = st[V];
    } else if (V - buf < len) {
      v = src[V - buf];
    }
    state_out[(int64_t)s * buf + i] = v;
  }
}

// numpy's float32 pairwise summation (numpy/_core/src/umath/loops_utils.h.src), which is what
// np.mean(pitch_values) executes at /root/reference/backend/services/prosody.py:90.
__device__ float np_pairwise_sum(const float* a, int64_t n) {
  if (n < 8) {
    float res = 0.f;
    for (int64_t i = 0; i < n; ++i) res = __fadd_rn(res, a[i]);
    return res;
  } else if (n <= 128) {
    float r[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = a[k];
    int64_t i;
    for (i = 8; i < n - (n % 8); i += 8) {
#pragma unroll
      for (int k = 0; k < 8; ++k) r[k] = __fadd_rn(r[k], a[i + k]);
    }
    float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                          __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __fadd_rn(res, a[i]);
    return res;
  } else {
    int64_t n2 = n / 2;
    n2 -= n2 % 8;
    return __fadd_rn(np_pairwise_sum(a, n2), np_pairwise_sum(a + n2, n - n2));
  }
}

// one warp per stream
__global__ void prosody_reduce_kernel(const float* __restrict__ f0, const float* __restrict__ hop_sumsq,
                                      const int64_t* __restrict__ hop_offsets,
                                      const int64_t* __restrict__ stream_lengths, int n_streams,
                                      float* __restrict__ scratch, float* __restrict__ rms_out,
                                      float* __restrict__ mean_f0_out, int32_t* __restrict__ n_voiced_out,
                                      uint8_t* __restrict__ tags_out) {
  const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (s >= n_streams) return;
  const int64_t h0 = hop_offsets[s];
  const int64_t nh = hop_offsets[s + 1] - h0;
  const int64_t len = stream_lengths[s];
  // whole-buffer RMS (prosody.py:67): fixed-order double accumulation of the per-hop sums
  double acc = 0.0;
  for (int64_t i = lane; i < nh; i += 32) acc += (double)hop_sumsq[h0 + i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, o);
  const float rms = len > 0 ? (float)sqrt(acc / (double)len) : nanf("");
  // ordered compaction of the voiced hops (prosody.py:86-87)
  int64_t cnt = 0;
  for (int64_t i0 = 0; i0 < nh; i0 += 32) {
    const int64_t i = i0 + lane;
    const float v = i < nh ? f0[h0 + i] : 0.f;
    const bool voiced = v > 0.f;
    const unsigned m = __ballot_sync(FULL_MASK, voiced);
    if (voiced) scratch[h0 + cnt + __popc(m & ((1u << lane) - 1u))] = v;
    cnt += __popc(m);
  }
  __syncwarp();
  if (lane == 0) {
    float mean = 0.f;
    uint8_t ptag = 1;  // 'Normal' when nothing is voiced (prosody.py:98-99)
    if (cnt > 0) {
      mean = __fdiv_rn(np_pairwise_sum(scratch + h0, cnt), (float)cnt);
      ptag = mean < 120.f ? 0 : (mean < 200.f ? 1 : 2);
    }
    // rms < 0.05 / 0.15 compared as numpy does for a float32 scalar
    const uint8_t etag = rms < 0.05f ? 0 : (rms < 0.15f ? 1 : 2);
    rms_out[s] = rms;
    mean_f0_out[s] = mean;
    n_voiced_out[s] = (int32_t)cnt;
    tags_out[2 * s + 0] = etag;
    tags_out[2 * s + 1] = ptag;
  }
}

// ------------------------------------------------------------------------------------------
// Per-chunk speech gate for many sessions: the step BEFORE the path.
// The reference asks Silero VAD per 1536-sample chunk (services/vad.py:40-77, engine.py:471-493); its model is a
// download and not on the path. This gate reuses what K4 already computes for the chunk's hops: a chunk is speech
// when its loudest hop exceeds level_db AND at least min_voiced of its hops are voiced (YIN found a period and
// the hop passed aubio's -50 dB silence gate). One thread per chunk; chunks are hops_per_chunk consecutive hops.
// ------------------------------------------------------------------------------------------
__global__ void vad_gate_kernel(const float* __restrict__ f0, const float* __restrict__ hop_db, long long n_chunks,
                                int hops_per_chunk, float level_db, int min_voiced, uint8_t* __restrict__ flags,
                                float* __restrict__ chunk_db) {
  const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_chunks) return;
  float mx = -INFINITY;
  int voiced = 0;
  for (int h = 0; h < hops_per_chunk; ++h) {
    const long long i = c * hops_per_chunk + h;
    mx = fmaxf(mx, hop_db[i]);
    voiced += f0[i] > 0.f;
  }
  flags[c] = (mx > level_db && voiced >= min_voiced) ? 1 : 0;
  if (chunk_db) chunk_db[c] = mx;
}

}  // namespace janus

using namespace janus;

extern "C" int32_t janus_yin_frames_per_group(void) { return YIN_F; }

extern "C" int32_t janus_yin_frames_f32(const float* pcm, const int64_t* stream_offsets,
                                        const int64_t* stream_lengths, const int64_t* hop_offsets,
                                        const int32_t* group_offsets, int32_t n_streams,
                                        int64_t total_groups, int32_t sample_rate, int32_t buf_size,
                                        int32_t hop_size, float tolerance, float silence_db,
                                        const float* state_in, float* f0_out, float* hop_db_out,
                                        float* hop_sumsq_out, void* stream) {
  JANUS_CHECK_ARG(n_streams >= 0 && total_groups >= 0, "negative sizes");
  if (n_streams == 0 || total_groups == 0) return JANUS_OK;
  JANUS_CHECK_ARG(pcm && stream_offsets && stream_lengths && hop_offsets && group_offsets,
                  "null input pointer");
  JANUS_CHECK_ARG(f0_out && hop_db_out && hop_sumsq_out, "null output pointer");
  JANUS_CHECK_ARG(buf_size >= 512 && buf_size % 512 == 0,
                  "buf_size %d unsupported: must be a positive multiple of 512", buf_size);
  JANUS_CHECK_ARG(hop_size > 0 && hop_size % 4 == 0 && hop_size <= buf_size,
                  "hop_size %d unsupported: must be a multiple of 4 and <= buf_size", hop_size);
  JANUS_CHECK_ARG(total_groups <= 0x7fffffffLL, "too many frame groups for one launch");
  YinParams p;
  p.pcm = pcm; p.stream_offsets = stream_offsets; p.stream_lengths = stream_lengths;
  p.hop_offsets = hop_offsets; p.group_offsets = group_offsets; p.state_in = state_in;
  p.f0_out = f0_out; p.hop_db_out = hop_db_out; p.hop_sumsq_out = hop_sumsq_out;
  p.n_streams = n_streams; p.sample_rate = sample_rate; p.buf = buf_size; p.hop = hop_size;
  const int NS = (YIN_F - 1) * hop_size + buf_size;
  p.ns_alloc = (NS + 16 + 63) / 64 * 64;
  p.tol = tolerance; p.silence_db = silence_db;
  const size_t smem = sizeof(float) * ((size_t)p.ns_alloc + (size_t)YIN_F * (buf_size / 2));
  JANUS_CHECK_ARG(smem <= 227 * 1024, "buf/hop combination needs %zu B of shared memory", smem);
  JANUS_CUDA(cudaFuncSetAttribute(yin_frames_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  yin_frames_kernel<<<(unsigned)total_groups, YIN_F * 32, smem, (cudaStream_t)stream>>>(p);
  JANUS_LAUNCH_CHECK("yin_frames_kernel");
  return JANUS_OK;
}

extern "C" int32_t janus_yin_state_f32(const float* pcm, const int64_t* stream_offsets,
                                       const int64_t* stream_lengths, int32_t n_streams,
                                       int32_t buf_size, int32_t hop_size, const float* state_in,
                                       float* state_out, void* stream) {
  if (n_streams <= 0) return JANUS_OK;
  JANUS_CHECK_ARG(pcm && stream_offsets && stream_lengths && state_out, "null pointer");
  JANUS_CHECK_ARG(state_in != state_out, "state_out must not alias state_in");
  JANUS_CHECK_ARG(buf_size > 0 && hop_size > 0, "bad buf/hop");
  yin_state_kernel<<<n_streams, 256, 0, (cudaStream_t)stream>>>(pcm, stream_offsets, stream_lengths,
                                                                buf_size, hop_size, state_in, state_out);
  JANUS_LAUNCH_CHECK("yin_state_kernel");
  return JANUS_OK;
}

extern "C" int32_t janus_prosody_reduce_f32(const float* f0, const float* hop_sumsq,
                                            const int64_t* hop_offsets, const int64_t* stream_lengths,
                                            int32_t n_streams, float* scratch, float* rms_out,
                                            float* mean_f0_out, int32_t* n_voiced_out,
                                            uint8_t* tags_out, void* stream) {
  if (n_streams <= 0) return JANUS_OK;
  JANUS_CHECK_ARG(f0 && hop_sumsq && hop_offsets && stream_lengths && scratch, "null input pointer");
  JANUS_CHECK_ARG(rms_out && mean_f0_out && n_voiced_out && tags_out, "null output pointer");
  const int wpb = 4;
  prosody_reduce_kernel<<<(n_streams + wpb - 1) / wpb, wpb * 32, 0, (cudaStream_t)stream>>>(
      f0, hop_sumsq, hop_offsets, stream_lengths, n_streams, scratch, rms_out, mean_f0_out,
      n_voiced_out, tags_out);
  JANUS_LAUNCH_CHECK("prosody_reduce_kernel");
  return JANUS_OK;
}

extern "C" int32_t janus_vad_gate_f32(const float* f0, const float* hop_db, int64_t n_chunks, int32_t hops_per_chunk,
                                      float level_db, int32_t min_voiced, uint8_t* flags_out, float* chunk_db_out,
                                      void* stream) {
  if (n_chunks <= 0) return JANUS_OK;
  JANUS_CHECK_ARG(f0 && hop_db && flags_out, "vad_gate: null pointer");
  JANUS_CHECK_ARG(hops_per_chunk >= 1 && min_voiced >= 0 && min_voiced <= hops_per_chunk, "vad_gate: bad hop counts");
  vad_gate_kernel<<<(unsigned)((n_chunks + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      f0, hop_db, n_chunks, hops_per_chunk, level_db, min_voiced, flags_out, chunk_db_out);
  JANUS_LAUNCH_CHECK("vad_gate_kernel");
  return JANUS_OK;
}

// ------------------------------------------------------------------------------------------
// No-allocation variant of janus_prosody_reduce_f32 (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops,
// harnesses that forbid it). The [total_hops] float scratch of the voiced-hop compaction lives in a file-scope
// device array of JANUS_PROSODY_STATIC_MAX_HOPS floats (4 MiB; every GPU that loads the library holds it until
// exit). In HOST code `g_prosody_scratch` is only a stand-in for the symbol, not a device address: the address is
// taken once per device, with that device current, through cudaGetSymbolAddress and passed to the kernel.
// Every call shares the array and it keeps the last call's contents; nothing has to start at zero (each stream's
// segment is written by the compaction before the pairwise sum reads it), so nothing is cleared. Two calls that
// overlap on the GPU (different streams, or a captured graph replayed concurrently with itself) race on it.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_prosody_scratch[JANUS_PROSODY_STATIC_MAX_HOPS];

extern "C" int32_t janus_prosody_reduce_static_f32(const float* f0, const float* hop_sumsq, const int64_t* hop_offsets,
                                                   const int64_t* stream_lengths, int32_t n_streams,
                                                   int64_t total_hops, float* rms_out, float* mean_f0_out,
                                                   int32_t* n_voiced_out, uint8_t* tags_out, void* stream) {
  if (n_streams <= 0) return JANUS_OK;
  JANUS_CHECK_ARG(f0 && hop_sumsq && hop_offsets && stream_lengths, "prosody_reduce_static: null input pointer");
  JANUS_CHECK_ARG(rms_out && mean_f0_out && n_voiced_out && tags_out, "prosody_reduce_static: null output pointer");
  JANUS_CHECK_ARG(total_hops >= 0, "prosody_reduce_static: negative total_hops");
  JANUS_CHECK_ARG(total_hops <= (int64_t)JANUS_PROSODY_STATIC_MAX_HOPS,
                  "prosody_reduce_static: total_hops=%lld exceeds the static scratch of %lld hops "
                  "(JANUS_PROSODY_STATIC_MAX_HOPS); use janus_prosody_reduce_f32 with caller-provided scratch",
                  (long long)total_hops, (long long)JANUS_PROSODY_STATIC_MAX_HOPS);
  static float* dev_ptr[64] = {};   // per device: the array has a different address on every GPU
  const int dev = current_device_slot();
  if (!dev_ptr[dev]) {
    float* ptr = nullptr;
    JANUS_CUDA(cudaGetSymbolAddress((void**)&ptr, g_prosody_scratch));
    dev_ptr[dev] = ptr;
  }
  const int wpb = 4;
  prosody_reduce_kernel<<<(n_streams + wpb - 1) / wpb, wpb * 32, 0, (cudaStream_t)stream>>>(
      f0, hop_sumsq, hop_offsets, stream_lengths, n_streams, dev_ptr[dev], rms_out, mean_f0_out, n_voiced_out, tags_out);
  JANUS_LAUNCH_CHECK("prosody_reduce_kernel (static scratch)");
  return JANUS_OK;
}
