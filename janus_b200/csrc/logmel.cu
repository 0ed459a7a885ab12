// K1: batched log-mel frontend (framed STFT -> power -> mel filterbank -> log10 -> clamp/scale).
//
// Replaces faster_whisper.feature_extractor.FeatureExtractor.__call__ + the per-window pad_or_trim
// that the reference reaches through /root/reference/backend/services/transcriber.py:53-57, and fuses
// the reference's own `audio_buffer[::3]` 48 kHz -> 16 kHz decimation (transcriber.py:51) into the
// sample loads (`decim` = 3).
//
// One warp = one 400-sample frame. The 400-point real DFT is a two-stage shared-memory mixed-radix
// transform, 400 = 20 x 20 (n = 20 n1 + n2, k = k1 + 20 k2):
//     Y[k1][n2] = sum_n1 x[20 n1 + n2] W20^(n1 k1)          k1 = 0..10 (real input: Y[20-k1] = conj Y[k1])
//     Z[k1][n2] = Y[k1][n2] W400^(n2 k1)
//     X[k1 + 20 k2] = sum_n2 Z[k1][n2] W20^(n2 k2)           only the 201 bins k <= 200
// i.e. ~25 k FMA per frame instead of 160 k for the direct DFT, all in fp32 (bf16/tf32 tensor-core
// DFTs do not hold the ~1e-5 parity of the log-mel values). The mel filterbank is applied in its
// sparse (triangular) form. Pass 1 writes log10(max(mel,1e-10)) and an utterance-global max
// (faster-whisper clamps to max-8 over the WHOLE utterance, not per window); pass 2 clamps, scales
// and emits the encoder's time-major bf16 layout (with the conv padding rows) and/or the reference's
// fp32 [n_mels, 3000] layout, zero-padding in FEATURE space past the utterance.
#include <cuda_bf16.h>
#include <math.h>

#include "common.cuh"

namespace janus {

constexpr int LM_NFFT = 400;
constexpr int LM_HOP = 160;
constexpr int LM_BINS = 201;
constexpr int LM_FRAMES = 3000;      // encoder window
constexpr int LM_TAIL_PAD = 160;     // FeatureExtractor(padding=160)
constexpr int LM_WARPS = 8;
constexpr int LM_ZLD = 21;         // row stride of Z in float2: 20 would put lanes k1 and k1 + 4 of stage 2 on the same banks
constexpr int LM_WARP_FLOATS = 400 + 2 * 20 * LM_ZLD + 208;   // per warp: xs[400] | Z[20][LM_ZLD] float2 | P[208]

__constant__ float2 c_w20[20];
__constant__ float2 c_w400[400];
__constant__ float c_hann[400];

static int init_tables() {
  static bool done[64] = {};
  int dev = 0;
  JANUS_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && done[dev]) return JANUS_OK;
  float2 w20[20], w400[400];
  float hann[400];
  const double PI = 3.14159265358979323846;
  for (int i = 0; i < 20; ++i) w20[i] = make_float2((float)cos(2 * PI * i / 20), (float)-sin(2 * PI * i / 20));
  for (int i = 0; i < 400; ++i) {
    w400[i] = make_float2((float)cos(2 * PI * i / 400), (float)-sin(2 * PI * i / 400));
    hann[i] = (float)(0.5 - 0.5 * cos(2 * PI * i / 400));  // np.hanning(401)[:-1]
  }
  JANUS_CUDA(cudaMemcpyToSymbol(c_w20, w20, sizeof(w20)));
  JANUS_CUDA(cudaMemcpyToSymbol(c_w400, w400, sizeof(w400)));
  JANUS_CUDA(cudaMemcpyToSymbol(c_hann, hann, sizeof(hann)));
  if (dev < 64) done[dev] = true;
  return JANUS_OK;
}

__device__ __forceinline__ int float_key(float f) {
  const int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

__global__ void __launch_bounds__(LM_WARPS * 32)
logmel_frames_kernel(const float* __restrict__ pcm, const int64_t* __restrict__ offsets,
                     const int64_t* __restrict__ lengths, int decim, int n_mels,
                     const int32_t* __restrict__ mel_start, const int32_t* __restrict__ mel_len,
                     const float* __restrict__ mel_w, int mel_max_len, float* __restrict__ scratch,
                     int* __restrict__ maxbuf, int nf_alloc, int nf_cap) {
  extern __shared__ __align__(16) float lm_smem[];
  float2* s_w20 = reinterpret_cast<float2*>(lm_smem);           // 20
  float2* s_w400 = s_w20 + 20;                                   // 400
  float* s_hann = reinterpret_cast<float*>(s_w400 + 400);        // 400
  float* s_melw = s_hann + 400;                                  // n_mels * mel_max_len
  int* s_mstart = reinterpret_cast<int*>(s_melw + n_mels * mel_max_len);
  int* s_mlen = s_mstart + n_mels;
  float* s_warp = reinterpret_cast<float*>(s_mlen + n_mels);     // per warp: LM_WARP_FLOATS
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 20; i += blockDim.x) s_w20[i] = c_w20[i];
  for (int i = threadIdx.x; i < 400; i += blockDim.x) { s_w400[i] = c_w400[i]; s_hann[i] = c_hann[i]; }
  for (int i = threadIdx.x; i < n_mels * mel_max_len; i += blockDim.x) s_melw[i] = mel_w[i];
  for (int i = threadIdx.x; i < n_mels; i += blockDim.x) { s_mstart[i] = mel_start[i]; s_mlen[i] = mel_len[i]; }
  __syncthreads();

  const int b = blockIdx.y;
  const int64_t len = (lengths[b] + decim - 1) / decim;  // samples after x[::decim]
  const int64_t L = len + LM_TAIL_PAD;
  int64_t nf = L / LM_HOP;                                // frames kept (last STFT frame dropped)
  if (nf > nf_cap) nf = nf_cap;
  const int t = blockIdx.x * LM_WARPS + warp;
  if (t >= nf) return;

  float* xs = s_warp + warp * LM_WARP_FLOATS;
  float2* Z = reinterpret_cast<float2*>(xs + 400);
  float* P = xs + 400 + 2 * 20 * LM_ZLD;
  const float* src = pcm + offsets[b];
  for (int n = lane; n < LM_NFFT; n += 32) {
    int64_t i = (int64_t)t * LM_HOP - LM_NFFT / 2 + n;
    if (i < 0) i = -i;                  // np.pad(mode="reflect")
    if (i >= L) i = 2 * (L - 1) - i;
    if (i < 0) i = 0;
    xs[n] = (i < len ? __ldg(src + i * decim) : 0.f) * s_hann[n];
  }
  __syncwarp();
  // stage 1 (+ twiddle): 11 x 20 outputs, mirrored into k1 = 11..19
  for (int o = lane; o < 220; o += 32) {
    const int k1 = o / 20, n2 = o - k1 * 20;
    float re = 0.f, im = 0.f;
    int tw = 0;
#pragma unroll
    for (int n1 = 0; n1 < 20; ++n1) {
      const float xv = xs[20 * n1 + n2];
      const float2 w = s_w20[tw];
      re = fmaf(xv, w.x, re);
      im = fmaf(xv, w.y, im);
      tw += k1;
      if (tw >= 20) tw -= 20;
    }
    Z[k1 * LM_ZLD + n2] = cmul(make_float2(re, im), s_w400[n2 * k1]);
    if (k1 >= 1 && k1 <= 9) Z[(20 - k1) * LM_ZLD + n2] = cmul(make_float2(re, -im), s_w400[n2 * (20 - k1)]);
  }
  __syncwarp();
  // stage 2: the 201 bins, power spectrum
  for (int k = lane; k < LM_BINS; k += 32) {
    const int k2 = k / 20, k1 = k - k2 * 20;
    float re = 0.f, im = 0.f;
    int tw = 0;
    const float2* z = Z + k1 * LM_ZLD;
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) {
      const float2 a = z[n2];
      const float2 w = s_w20[tw];
      re += a.x * w.x - a.y * w.y;
      im += a.x * w.y + a.y * w.x;
      tw += k2;
      if (tw >= 20) tw -= 20;
    }
    P[k] = re * re + im * im;
  }
  __syncwarp();
  float mx = -INFINITY;
  float* orow = scratch + ((int64_t)b * nf_alloc + t) * n_mels;
  for (int m = lane; m < n_mels; m += 32) {
    const int st = s_mstart[m], ln = s_mlen[m];
    const float* w = s_melw + m * mel_max_len;
    float acc = 0.f;
    for (int j = 0; j < ln; ++j) acc = fmaf(w[j], P[st + j], acc);
    const float lg = log10f(fmaxf(acc, 1e-10f));
    orow[m] = lg;
    mx = fmaxf(mx, lg);
  }
  mx = warp_max(mx);
  if (lane == 0) atomicMax(maxbuf + b, float_key(mx));
}

// Window w = frames [seek, seek + 3000) of utterance u = win_utt[w] (win_utt == NULL: u = w, seek = 0): the slice +
// pad_or_trim of faster-whisper's transcribe() loop, with the clamp floor of the WHOLE utterance.
__global__ void logmel_finalize_kernel(const float* __restrict__ scratch, const int* __restrict__ maxbuf,
                                       const int64_t* __restrict__ lengths, int decim, int n_mels, int nf_alloc,
                                       int nf_cap, const int32_t* __restrict__ win_utt,
                                       const int32_t* __restrict__ win_seek,
                                       __nv_bfloat16* __restrict__ out_bf16, float* __restrict__ out_f32) {
  const int b = blockIdx.y;
  const int u = win_utt ? win_utt[b] : b;
  const int seek = win_utt ? win_seek[b] : 0;
  const int64_t len = (lengths[u] + decim - 1) / decim;
  int64_t nf = (len + LM_TAIL_PAD) / LM_HOP;
  if (nf > nf_cap) nf = nf_cap;
  const int seg = (int)max((int64_t)0, min((int64_t)LM_FRAMES, nf - 1 - seek));  // content frames of this window
  const float floor_v = key_float(maxbuf[u]) - 8.0f;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // over (LM_FRAMES + 2) * n_mels, time-major
  if (idx >= (LM_FRAMES + 2) * n_mels) return;
  const int row = idx / n_mels, m = idx - row * n_mels;
  const int t = row - 1;
  float y = 0.f;
  if (t >= 0 && t < seg) {
    const float v = scratch[((int64_t)u * nf_alloc + seek + t) * n_mels + m];
    y = (fmaxf(v, floor_v) + 4.0f) / 4.0f;
  }
  if (out_bf16) out_bf16[((int64_t)b * (LM_FRAMES + 2) + row) * n_mels + m] = __float2bfloat16_rn(y);
  if (out_f32 && t >= 0 && t < LM_FRAMES) out_f32[((int64_t)b * n_mels + m) * LM_FRAMES + t] = y;
}

}  // namespace janus

using namespace janus;

extern "C" int32_t janus_logmel_frames_alloc(void) { return LM_FRAMES + 1; }

static int32_t logmel_frames(const float* pcm, const int64_t* offsets, const int64_t* lengths, int32_t n, int32_t decim,
                             int32_t n_mels, const int32_t* mel_start, const int32_t* mel_len, const float* mel_w,
                             int32_t mel_max_len, float* scratch, int32_t* maxbuf, int32_t nf_alloc, cudaStream_t st) {
  JANUS_CHECK_ARG(pcm && offsets && lengths && mel_start && mel_len && mel_w && scratch && maxbuf, "logmel: null pointer");
  JANUS_CHECK_ARG(decim >= 1 && n_mels > 0 && n_mels <= 128 && mel_max_len > 0 && mel_max_len <= 64,
                  "logmel: unsupported decim/n_mels/mel_max_len");
  JANUS_CHECK_ARG(nf_alloc >= 1, "logmel: nf_alloc must be >= 1");
  int rc = init_tables();
  if (rc) return rc;
  JANUS_CUDA(cudaMemsetAsync(maxbuf, 0x80, sizeof(int32_t) * n, st));
  const size_t smem = sizeof(float2) * 420 + sizeof(float) * 400 + sizeof(float) * n_mels * mel_max_len +
                      sizeof(int) * 2 * n_mels + sizeof(float) * LM_WARPS * LM_WARP_FLOATS;
  JANUS_CUDA(cudaFuncSetAttribute(logmel_frames_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const dim3 grid((nf_alloc + LM_WARPS - 1) / LM_WARPS, n);
  logmel_frames_kernel<<<grid, LM_WARPS * 32, smem, st>>>(pcm, offsets, lengths, decim, n_mels, mel_start, mel_len,
                                                          mel_w, mel_max_len, scratch, maxbuf, nf_alloc, nf_alloc);
  JANUS_LAUNCH_CHECK("logmel_frames_kernel");
  return JANUS_OK;
}

extern "C" int32_t janus_logmel_f32(const float* pcm, const int64_t* offsets, const int64_t* lengths, int32_t n,
                                    int32_t decim, int32_t n_mels, const int32_t* mel_start,
                                    const int32_t* mel_len, const float* mel_w, int32_t mel_max_len,
                                    float* scratch, int32_t* maxbuf, void* mel_bf16_out, float* mel_f32_out,
                                    void* stream) {
  if (n <= 0) return JANUS_OK;
  JANUS_CHECK_ARG(mel_bf16_out || mel_f32_out, "logmel: no output requested");
  cudaStream_t st = (cudaStream_t)stream;
  const int nf_alloc = LM_FRAMES + 1;
  int rc = logmel_frames(pcm, offsets, lengths, n, decim, n_mels, mel_start, mel_len, mel_w, mel_max_len, scratch,
                         maxbuf, nf_alloc, st);
  if (rc) return rc;
  const int total = (LM_FRAMES + 2) * n_mels;
  const dim3 grid2((total + 255) / 256, n);
  logmel_finalize_kernel<<<grid2, 256, 0, st>>>(scratch, maxbuf, lengths, decim, n_mels, nf_alloc, nf_alloc, nullptr,
                                                nullptr, static_cast<__nv_bfloat16*>(mel_bf16_out), mel_f32_out);
  JANUS_LAUNCH_CHECK("logmel_finalize_kernel");
  return JANUS_OK;
}

extern "C" int32_t janus_logmel_utterance_f32(const float* pcm, const int64_t* offsets, const int64_t* lengths,
                                              int32_t n, int32_t decim, int32_t n_mels, const int32_t* mel_start,
                                              const int32_t* mel_len, const float* mel_w, int32_t mel_max_len,
                                              float* scratch, int32_t nf_alloc, int32_t* maxbuf, void* stream) {
  if (n <= 0) return JANUS_OK;
  return logmel_frames(pcm, offsets, lengths, n, decim, n_mels, mel_start, mel_len, mel_w, mel_max_len, scratch,
                       maxbuf, nf_alloc, (cudaStream_t)stream);
}

extern "C" int32_t janus_logmel_windows_f32(const float* scratch, const int32_t* maxbuf, const int64_t* lengths,
                                            int32_t decim, int32_t n_mels, int32_t nf_alloc, const int32_t* win_utt,
                                            const int32_t* win_seek, int32_t n_windows, void* mel_bf16_out,
                                            float* mel_f32_out, void* stream) {
  if (n_windows <= 0) return JANUS_OK;
  JANUS_CHECK_ARG(scratch && maxbuf && lengths && win_utt && win_seek, "logmel_windows: null pointer");
  JANUS_CHECK_ARG(mel_bf16_out || mel_f32_out, "logmel_windows: no output requested");
  JANUS_CHECK_ARG(decim >= 1 && n_mels > 0 && n_mels <= 128 && nf_alloc >= 1, "logmel_windows: bad arguments");
  const int total = (LM_FRAMES + 2) * n_mels;
  const dim3 grid2((total + 255) / 256, n_windows);
  logmel_finalize_kernel<<<grid2, 256, 0, (cudaStream_t)stream>>>(
      scratch, maxbuf, lengths, decim, n_mels, nf_alloc, nf_alloc, win_utt, win_seek,
      static_cast<__nv_bfloat16*>(mel_bf16_out), mel_f32_out);
  JANUS_LAUNCH_CHECK("logmel_finalize_kernel");
  return JANUS_OK;
}
