/*
 * janus_b200.h -- C ABI of libjanus_b200.so (hand-written sm_100a CUDA kernels for the
 * sender-side semantic-encode path of akshatvasisht/janus).
 *
 * Conventions
 *   - every entry point is extern "C", returns int32 status: 0 = OK, <0 = janus_status_t error;
 *     janus_last_error() returns a thread-local, human-readable message for the last failure.
 *   - all tensor arguments are DEVICE pointers owned by the caller (no torch types); shapes are
 *     explicit; `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - the library allocates nothing persistent and never synchronises the host, except for the
 *     *_host convenience entry points which say so.
 *   - native code never calls abort()/exit(); the Python service layer turns a non-zero status
 *     into an ordinary Python exception, which is the error contract the reference's callers
 *     rely on (/root/reference/backend/services/engine.py:513-525).
 *
 * What each entry point replaces in the reference (paths relative to /root/reference):
 *   janus_yin_frames_f32      aubio.pitch('yin',4096,hop,sr) called once per hop
 *                             backend/services/prosody.py:32-34 (ctor), :76-87 (hop loop);
 *                             upstream aubio 0.4.9 src/pitch/pitch.c + pitchyin.c + mathutils.c
 *   janus_prosody_reduce_f32  backend/services/prosody.py:67 (RMS), :69-74 (energy tag),
 *                             :86-99 (voiced filter, np.mean, pitch tag)
 *   janus_yin_state_f32       the stateful sliding buffer inside the aubio object
 *                             (prosody.py:32; aubio_pitch_slideblock)
 */
#ifndef JANUS_B200_H
#define JANUS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define JANUS_ABI_VERSION 2   /* 2: janus_set_option replaces janus_set_pdl; debug exports and the fused-segment flag retired */

typedef enum {
  JANUS_OK = 0,
  JANUS_ERR_INVALID_ARG = -1,  /* bad shape / unsupported parameter combination */
  JANUS_ERR_CUDA = -2,         /* a CUDA runtime call or launch failed           */
  JANUS_ERR_UNSUPPORTED = -3,  /* device is not sm_100 / feature not built       */
  JANUS_ERR_WORKSPACE = -4     /* caller-provided workspace too small            */
} janus_status_t;

int32_t janus_abi_version(void);
const char* janus_last_error(void);
/* 0 when device `dev` can run the kernels (compute capability 10.x), else JANUS_ERR_UNSUPPORTED. */
int32_t janus_check_device(int32_t dev);
/* Process-wide tunables, set explicitly (nothing in the library reads the environment). value < 0 restores the
 * built-in default. They select between measured alternates of a kernel, for A/B measurements and so that both
 * variants can be tested in one process; results are identical within the stated tolerances either way. */
typedef enum {
  JANUS_OPT_PDL = 0,                   /* programmatic dependent launch for the decode-step chain (default 1)       */
  JANUS_OPT_ATT_SPLIT = 1,             /* encoder attention: softmax threads per query row, 1 | 2 (default 1)        */
  JANUS_OPT_XLAT_STREAM = 2,           /* latent cross-attention: 1 = streaming two-pass kernel even for d <= 768    */
  JANUS_OPT_XLAT_CTAS = 3,             /* latent cross-attention: cap of the persistent grid, 0 = one CTA per SM     */
  JANUS_OPT_XA_CTAS_PER_SM = 4,        /* K/V cross-attention: cap of CTAs per SM, 0 = no cap                        */
  JANUS_OPT_BEAM_SHARED_SELF_ATTN = 5, /* beam search: shared-prefix self-attention kernel (default 1) or per row    */
  JANUS_OPT_COUNT = 6
} janus_option_t;
int32_t janus_set_option(int32_t option, int32_t value);

/* ------------------------------------------------------------------------------------------
 * K4: YIN pitch + per-hop level (bit-exact restatement of aubio's float32 arithmetic)
 * ------------------------------------------------------------------------------------------
 * Streams are concatenated in `pcm`; stream s occupies pcm[stream_offsets[s] .. +stream_lengths[s]).
 * n_hops(s) = ceil(len/hop); hop_offsets[s] = sum_{t<s} n_hops(t) (n_streams+1 entries).
 * Frames of a stream are processed in groups of janus_yin_frames_per_group() consecutive hops
 * by one thread block; group_offsets[s] = sum_{t<s} ceil(n_hops(t)/F) (n_streams+1 entries).
 * state_in: [n_streams, buf_size] previous sliding buffers, or NULL for fresh (all-zero) state.
 * Outputs (all [total_hops]): f0_out in Hz (0 = unvoiced / gated), hop_db_out = 10*log10(mean(x^2))
 * of the hop (aubio_db_spl), hop_sumsq_out = sum(x^2) of the hop (for the stream RMS).
 * Constraints: buf_size % 512 == 0, hop_size % 4 == 0, hop_size <= buf_size.
 */
int32_t janus_yin_frames_per_group(void);
int32_t janus_yin_frames_f32(const float* pcm, const int64_t* stream_offsets,
                             const int64_t* stream_lengths, const int64_t* hop_offsets,
                             const int32_t* group_offsets, int32_t n_streams, int64_t total_groups,
                             int32_t sample_rate, int32_t buf_size, int32_t hop_size,
                             float tolerance, float silence_db, const float* state_in,
                             float* f0_out, float* hop_db_out, float* hop_sumsq_out, void* stream);

/* New sliding-buffer state after the call above: state_out[s] = last buf_size samples of
 * (state_in[s] ++ zero-padded stream s). state_out must not alias state_in. */
int32_t janus_yin_state_f32(const float* pcm, const int64_t* stream_offsets,
                            const int64_t* stream_lengths, int32_t n_streams, int32_t buf_size,
                            int32_t hop_size, const float* state_in, float* state_out, void* stream);

/* Per-stream reduction: rms_out[s] = sqrt(sum(x^2)/len); mean_f0_out[s] = float32 mean of the
 * voiced (f0 > 0) hops in numpy's pairwise summation order (bit-identical to np.mean of the
 * float32 list at prosody.py:90), 0 when no hop is voiced; n_voiced_out[s]; tags_out[2*s+0] =
 * energy tag (0 Quiet, 1 Normal, 2 Loud), tags_out[2*s+1] = pitch tag (0 Deep, 1 Normal, 2 High).
 * scratch: [total_hops] floats. */
int32_t janus_prosody_reduce_f32(const float* f0, const float* hop_sumsq, const int64_t* hop_offsets,
                                 const int64_t* stream_lengths, int32_t n_streams, float* scratch,
                                 float* rms_out, float* mean_f0_out, int32_t* n_voiced_out,
                                 uint8_t* tags_out, void* stream);

/* janus_prosody_reduce_f32 without caller-provided scratch and WITHOUT ANY ALLOCATION. This entry point has no
 * counterpart in the reference: it was added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The scratch of the
 * voiced-hop compaction is a static device array of JANUS_PROSODY_STATIC_MAX_HOPS floats (4 MiB, held by every GPU
 * that loads the library until exit). total_hops must be hop_offsets[n_streams]; a value beyond the maximum is
 * rejected with JANUS_ERR_INVALID_ARG and a message. Results are bit-identical to janus_prosody_reduce_f32. All calls
 * share the array (it keeps the previous call's contents; nothing needs clearing): calls that OVERLAP on the GPU --
 * two streams, or a replayed graph running concurrently with another call -- race. The first call on a device takes
 * the array's address (cudaGetSymbolAddress): make one call outside stream capture before capturing it. Calls no
 * cudaMalloc / cudaFree / allocating library routine. */
#define JANUS_PROSODY_STATIC_MAX_HOPS 1048576
int32_t janus_prosody_reduce_static_f32(const float* f0, const float* hop_sumsq, const int64_t* hop_offsets,
                                        const int64_t* stream_lengths, int32_t n_streams, int64_t total_hops,
                                        float* rms_out, float* mean_f0_out, int32_t* n_voiced_out, uint8_t* tags_out,
                                        void* stream);

/* Per-chunk speech gate for many sessions, from the outputs of janus_yin_frames_f32 (the step before the path:
 * backend/services/engine.py:471-493 asks backend/services/vad.py per 1536-sample chunk; Silero's model is a download,
 * so this is an energy gate (two thresholds), NOT a reproduction of Silero's decisions). Chunk c = hops
 * [c*hops_per_chunk, (c+1)*hops_per_chunk) of f0 / hop_db; flags_out[c] = 1 iff max hop level > level_db and at
 * least min_voiced hops have f0 > 0 (passed aubio's -50 dB silence gate). chunk_db_out (or NULL) receives the max hop level. */
int32_t janus_vad_gate_f32(const float* f0, const float* hop_db, int64_t n_chunks, int32_t hops_per_chunk,
                           float level_db, int32_t min_voiced, uint8_t* flags_out, float* chunk_db_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * K2 building block: bf16 GEMM on tcgen05/TMEM fed by TMA, fused epilogue
 * ------------------------------------------------------------------------------------------
 * out[b, t + out_row_offset, n] = act( sum_k A[b, t, k] * W[n, k] + bias[n] ) + pos[t, n] + residual[b, t, n]
 * A: bf16, row t of batch b at A + b*a_batch_stride + t*lda (elements; rows may overlap: this is how
 * the conv1d stem layers are expressed), W: bf16 [N, K] with leading dimension ldw, out/residual bf16
 * with leading dimension ldc and batch stride out_batch_stride, bias fp32 [N] or NULL, pos fp32
 * [M_per_batch, N] or NULL, act 0 = none / 1 = exact (erf) GELU. N % 64 == 0; K, lda, ldw, ldc % 8 == 0.
 * Replaces the int8/fp32 CPU GEMMs inside CTranslate2's Whisper.encode that the reference reaches via
 * backend/services/transcriber.py:53-57. */
int32_t janus_gemm_bf16(const void* A, int64_t lda, int64_t a_batch_stride, const void* W, int64_t ldw,
                        void* out, int64_t ldc, int64_t out_batch_stride, int32_t out_row_offset,
                        int32_t M_per_batch, int32_t n_batch, int32_t N, int32_t K, const float* bias,
                        const void* residual, const float* pos, int32_t act, void* stream);

/* Building blocks exposed for parity tests (same kernels the encoder driver launches). */
int32_t janus_layernorm_bf16(const void* x, const float* w, const float* b, void* y, int64_t rows, int32_t d,
                             void* stream);
/* Fused non-causal attention over packed qkv bf16 [n_batch, T, 3*64*n_heads] -> out bf16 [n_batch, T, 64*n_heads]. */
int32_t janus_attention_enc_bf16(const void* qkv, void* out, int32_t n_batch, int32_t T, int32_t n_heads,
                                 void* stream);

/* ------------------------------------------------------------------------------------------
 * K1: batched log-mel frontend (replaces faster_whisper FeatureExtractor + pad_or_trim; fuses the
 * reference's audio[::3] decimation, backend/services/transcriber.py:51, as decim = 3)
 * ------------------------------------------------------------------------------------------
 * Utterance s = pcm[offsets[s] .. +lengths[s]) at the SOURCE rate; the frontend sees x[::decim].
 * Utterances longer than one 30 s window are clipped to the first window (the host splits them).
 * mel_start/mel_len/mel_w: the Slaney filterbank in sparse row form (row m covers bins
 * [mel_start[m], mel_start[m]+mel_len[m]) with weights mel_w[m*mel_max_len + j]).
 * scratch: fp32 [n, janus_logmel_frames_alloc(), n_mels]; maxbuf: int32 [n].
 * mel_bf16_out (or NULL): bf16 [n, 3002, n_mels] time-major with zero rows 0 and 3001 (the encoder's
 * conv padding); mel_f32_out (or NULL): fp32 [n, n_mels, 3000] (the reference's layout). */
int32_t janus_logmel_frames_alloc(void);
int32_t janus_logmel_f32(const float* pcm, const int64_t* offsets, const int64_t* lengths, int32_t n,
                         int32_t decim, int32_t n_mels, const int32_t* mel_start, const int32_t* mel_len,
                         const float* mel_w, int32_t mel_max_len, float* scratch, int32_t* maxbuf,
                         void* mel_bf16_out, float* mel_f32_out, void* stream);

/* The same in two calls, for utterances longer than one 30 s window (faster-whisper computes the features ONCE per
 * utterance -- one global max-8 clamp floor, a continuous STFT -- and then slices 3000-frame windows at `seek` frames,
 * advancing seek by the last decoded timestamp; the call site is backend/services/transcriber.py:53-57):
 *   janus_logmel_utterance_f32  pass 1 over whole utterances: scratch fp32 [n, nf_alloc, n_mels] = log10 mel of up
 *                               to nf_alloc frames of each utterance (frame count of utterance s:
 *                               (ceil(len/decim) + 160) / 160), maxbuf int32 [n] = ordered-int key of its maximum;
 *   janus_logmel_windows_f32    pass 2 for n_windows windows: window w = frames [win_seek[w], +3000) of utterance
 *                               win_utt[w], clamped to that utterance's max - 8, zero-padded in feature space. */
int32_t janus_logmel_utterance_f32(const float* pcm, const int64_t* offsets, const int64_t* lengths, int32_t n,
                                   int32_t decim, int32_t n_mels, const int32_t* mel_start, const int32_t* mel_len,
                                   const float* mel_w, int32_t mel_max_len, float* scratch, int32_t nf_alloc,
                                   int32_t* maxbuf, void* stream);
int32_t janus_logmel_windows_f32(const float* scratch, const int32_t* maxbuf, const int64_t* lengths, int32_t decim,
                                 int32_t n_mels, int32_t nf_alloc, const int32_t* win_utt, const int32_t* win_seek,
                                 int32_t n_windows, void* mel_bf16_out, float* mel_f32_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * K2: Whisper encoder (replaces CTranslate2 Whisper.encode reached from transcriber.py:53-57)
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  int32_t n_mels, n_ctx, d, n_heads, n_layers;   /* n_ctx = 1500; input has 2*n_ctx frames */
} janus_whisper_dims_t;

typedef struct {          /* device pointers; matrices bf16 row-major [out, in], vectors fp32 */
  const float* ln1_w; const float* ln1_b;
  const void* w_qkv;  const float* b_qkv;    /* [3d, d]; the key third of b_qkv is zero (Whisper has no key bias) */
  const void* w_o;    const float* b_o;
  const float* ln2_w; const float* ln2_b;
  const void* w_fc1;  const float* b_fc1;    /* [4d, d] */
  const void* w_fc2;  const float* b_fc2;    /* [d, 4d] */
} janus_enc_layer_t;

typedef struct {
  const void* conv1_w; const float* conv1_b;  /* bf16 [d, 3*n_mels], tap-major: [d][k][c] */
  const void* conv2_w; const float* conv2_b;  /* bf16 [d, 3*d],      tap-major */
  const float* pos;                            /* fp32 [n_ctx, d] sinusoids */
  const float* ln_post_w; const float* ln_post_b;
  const janus_enc_layer_t* layers;             /* HOST array, n_layers entries */
} janus_encoder_weights_t;

int64_t janus_whisper_encode_workspace_bytes(const janus_whisper_dims_t* dims, int32_t batch);
/* mel_pad: bf16 [batch, 2*n_ctx+2, n_mels] (janus_logmel_f32's bf16 output); enc_out: bf16 [batch, n_ctx, d]. */
int32_t janus_whisper_encode_bf16(const janus_whisper_dims_t* dims, const janus_encoder_weights_t* weights,
                                  const void* mel_pad, int32_t batch, void* enc_out, void* workspace,
                                  int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * K3: Whisper decoder, greedy / temperature 0 (replaces CTranslate2 Whisper.generate + the
 * faster-whisper logit rules executed while transcriber.py:59-64 iterates `segments`)
 * ------------------------------------------------------------------------------------------ */
typedef struct {          /* device pointers; matrices bf16 row-major [out, in], vectors fp32 */
  const float* ln1_w; const float* ln1_b;
  const void* w_qkv;  const float* b_qkv;    /* self-attention [3d, d]; key bias third is zero */
  const void* w_o;    const float* b_o;
  const float* lnc_w; const float* lnc_b;
  const void* w_cq;   const float* b_cq;     /* cross-attention query / output projections */
  const void* w_co;   const float* b_co;
  const float* ln2_w; const float* ln2_b;
  const void* w_fc1;  const float* b_fc1;
  const void* w_fc2;  const float* b_fc2;
  /* JANUS_DECODE_LATENT only (else NULL): the layer's cross-attention key projection regrouped per head,
   * bf16 [H*d, 64] with row h*d + n = Wk[h*64 .. h*64+63][n] (so that q'_h = Wk_h^T q_h is a grouped GEMM),
   * and its value projection bf16 [d, d] / fp32 [d] as separate per-layer tensors */
  const void* w_ck_lat;
  const void* w_cv;   const float* b_cv;
} janus_dec_layer_t;

typedef struct {
  const void* tok_emb;                 /* bf16 [n_vocab_pad, d], rows >= n_vocab are zero (tied output head) */
  const float* pos;                    /* fp32 [n_text_ctx, d] learned positions */
  const float* ln_w; const float* ln_b;
  const void* w_ckv; const float* b_ckv; /* all layers' cross K/V projections stacked: bf16 [L*2*d, d] as
                                            [l][k|v][d]; fp32 bias with zeros in the key parts */
  const uint8_t* suppress_mask;        /* [n_vocab] 1 = never emit (non-speech set + specials) */
  const janus_dec_layer_t* layers;     /* HOST array, n_layers entries */
  int32_t n_vocab, n_vocab_pad, n_text_ctx, kv_ld;   /* kv_ld = padded key count of the transposed cross-K cache */
  int32_t eot, blank, timestamp_begin, max_initial_timestamp_index;
} janus_decoder_weights_t;

/* JANUS_DECODE_LIGHT: run every linear layer of the step through the register-streaming kernel (no
 * TMEM, < 32 KB shared memory) so that the step can share SMs with the tcgen05 encoder kernels of
 * another stream; requires batch <= 64. */
#define JANUS_DECODE_LIGHT 1
/* (flag value 2 is retired: the experimental grid-barrier segment kernel was measured slower than the per-op chain
 * and is no longer part of the library) */
/* JANUS_DECODE_LATENT: cross-attention reads the encoder states (state->enc_out) instead of per-layer K/V
 * caches -- half the HBM bytes per step, no cross_k / cross_v at all (janus_b200/csrc/dec_xlat.cu). Needs
 * d % 128 == 0, d <= 1024, <= 16 heads, the w_ck_lat / w_cv / b_cv layer weights and state->enc_out / qlat / zlat. */
#define JANUS_DECODE_LATENT 4
/* JANUS_DECODE_KV_SWAPPED (beam search, K/V-cache form): the two cross-attention cache layouts swap roles --
 * cross_k bf16 [L, n_streams, H, n_ctx, 64] (row-major) and cross_v bf16 [L, n_streams, H, 64, kv_ld] (positions
 * contiguous), as written by janus_whisper_cross_kv_swapped_bf16 -- so that both contractions of the decode-time
 * cross-attention are mma.sync tiles fed by plain 16-byte global loads (janus_b200/csrc/decoder.cu). */
#define JANUS_DECODE_KV_SWAPPED 8

typedef struct {
  int32_t batch, max_len, n_prompt;    /* max_len = n_prompt + max_new_tokens */
  int32_t flags;                       /* JANUS_DECODE_* */
  int32_t self_ld;                     /* positions allocated per (stream, head) in self_k / self_v; 0 = n_text_ctx.
                                          Must be >= max_len (a session that decodes 131 tokens need not hold 448) */
  int32_t reserved0;
  int32_t* tokens;                     /* [batch, max_len]; the sot prompt is pre-filled by the caller */
  const int32_t* forced;               /* [batch, max_len] teacher forcing (-1 = free) or NULL */
  int32_t* choice;                     /* [batch, max_len] argmax under the rules (diagnostic) or NULL */
  float* logprob;                      /* [batch, max_len] log-prob of the rule-filtered argmax */
  float* sum_logprob;                  /* [batch] */
  int32_t* pos;                        /* device scalar: index of the token the next step consumes */
  uint8_t* done;                       /* [batch] */
  int32_t* n_done;                     /* device scalar */
  void* self_k; void* self_v;          /* bf16 [L, batch, H, self_ld, 64] */
  const void* cross_k;                 /* bf16 [L, batch, H, 64, kv_ld]  (keys contiguous); NULL with JANUS_DECODE_LATENT */
  const void* cross_v;                 /* bf16 [L, batch, H, n_ctx, 64];                   NULL with JANUS_DECODE_LATENT */
  void* x; void* xn; void* qkv; void* att; void* q; void* ff;  /* bf16 [batch, d|d|3d|d|d|4d] */
  float* logits;                       /* fp32 [batch, n_vocab_pad] */
  uint64_t* reserved1;                 /* NULL (was the grid-barrier counter of the retired fused segment kernel) */
  const void* enc_out;                 /* JANUS_DECODE_LATENT: bf16 [batch, n_ctx, d] encoder states, else NULL */
  void* qlat;                          /* JANUS_DECODE_LATENT: bf16 [batch, 16, d], zero-initialised once */
  void* zlat;                          /* JANUS_DECODE_LATENT: bf16 [batch, H, d] */
} janus_decode_state_t;

/* Skinny linear layer of the decode step: out[r, n] = act(sum_k A[r,k] W[n,k] + bias[n]) (+ residual[r,n]),
 * rows <= 256, A bf16 [rows, K] (leading dim lda), W bf16 [N, K], out bf16 or fp32 [rows, N]. */
int32_t janus_dec_linear_bf16(const void* A, int32_t lda, const void* W, const float* bias, const void* residual,
                              void* out, int32_t rows, int32_t N, int32_t K, int32_t act, int32_t out_f32,
                              void* stream);
/* One layer of decode-time cross-attention (the HBM-bound kernel of the step), exposed for parity tests
 * and roofline timing: q bf16 [batch, 64*n_heads]; cross_k bf16 [batch, n_heads, 64, kv_ld];
 * cross_v bf16 [batch, n_heads, n_ctx, 64]; out bf16 [batch, 64*n_heads]. */
int32_t janus_dec_cross_attn_bf16(const void* q, const void* cross_k, const void* cross_v, void* out,
                                  int32_t batch, int32_t n_heads, int32_t n_ctx, int32_t kv_ld, void* stream);
/* Decode-time cross-attention over the encoder states themselves ("latent" form, K/V projections absorbed
 * into the query and output sides -- janus_b200/csrc/dec_xlat.cu): qlat bf16 [n_windows, 16, d] (absorbed
 * queries Wk_h^T q_h, rows >= n_heads zero), enc bf16 [n_windows, n_ctx, d]; z bf16 [n_windows, n_heads, d]
 * = softmax_t(enc . qlat_h / 8) @ enc. Needs d % 128 == 0, d <= 1024, d == 64 * n_heads, n_heads <= 16
 * (d <= 768: one-tile ring kernel; d = 1024: streaming two-pass kernel). */
int32_t janus_dec_xlat_bf16(const void* qlat, const void* enc, void* z, int32_t n_windows, int32_t n_ctx, int32_t d,
                            int32_t n_heads, void* stream);
/* Cross-attention K/V of all decoder layers for a batch of encoder outputs (one tcgen05 GEMM whose
 * epilogue writes the two cache layouts directly). dims->n_layers/d/n_heads describe the DECODER. */
int32_t janus_whisper_cross_kv_bf16(const janus_whisper_dims_t* dims, const janus_decoder_weights_t* weights,
                                    const void* enc_out, int32_t batch, void* cross_k, void* cross_v, void* stream);
/* One decode step for every stream of the batch: consumes tokens[:, *pos], emits tokens[:, *pos+1],
 * then *pos += 1. Nothing in it depends on host-side scalars, so it can be captured once in a CUDA
 * graph and replayed. */
int32_t janus_whisper_decode_step(const janus_whisper_dims_t* dims, const janus_decoder_weights_t* weights,
                                  const janus_decode_state_t* state, void* stream);

/* ------------------------------------------------------------------------------------------
 * K3, beam search (BASELINE configs[3]: beam 5). Replaces what CTranslate2 `Whisper.generate(beam_size=5,
 * patience=1, length_penalty=1)` does when faster-whisper's `transcribe(beam_size=5)` is called; the reference's
 * own call site passes beam_size=1 (backend/services/transcriber.py:55), which is janus_whisper_decode_step.
 * Algorithm = the published OpenAI Whisper BeamSearchDecoder / MaximumLikelihoodRanker (oracle/whisper_ref.py
 * beam_decode states it step by step).
 *
 * Rows of the decode state are the beams of n_streams streams: state->batch == n_streams * beam, row = stream * beam
 * + k; state->tokens is [rows, max_len] with the prompt pre-filled in every row. Cross-attention inputs exist once
 * per STREAM (state->enc_out [n_streams, n_ctx, d] or cross_k / cross_v with batch = n_streams); the self-attention
 * caches hold one slot per row and are never copied on a beam reorder: indir[row][j] names the slot (0 .. beam-1
 * within the stream) that holds position j of row's hypothesis. state->forced / choice / logprob / sum_logprob /
 * done are not used.
 * ------------------------------------------------------------------------------------------ */
#define JANUS_MAX_BEAM 8
typedef struct {
  int32_t beam, n_streams;
  int32_t max_candidates;              /* round(beam * patience): finished hypotheses kept per stream */
  int32_t reserved0;
  float* cum;                          /* [rows] cumulative log-prob of the live beams; reset to 0, -inf, -inf, ... per stream */
  uint8_t* indir;                      /* [rows, self_ld]; reset to indir[row][j] = row % beam */
  float* cand_score;                   /* [rows, beam + 1] scratch: candidates of the current step */
  int32_t* cand_tok;                   /* [rows, beam + 1] */
  int32_t* fin_tokens;                 /* [n_streams, max_candidates, max_len] finished hypotheses (prompt .. eot) */
  float* fin_score;                    /* [n_streams, max_candidates] their sum of log-probs (eot included) */
  int32_t* fin_len;                    /* [n_streams, max_candidates] sampled tokens before eot */
  int32_t* n_fin;                      /* [n_streams]; reset to 0 */
  uint8_t* stream_done;                /* [n_streams]; 1 once n_fin == max_candidates (state->n_done counts them) */
} janus_beam_state_t;

/* One beam-search step for every stream: consumes tokens[:, *pos] of every row, chooses the next beams, rewrites the
 * rows' token histories and the indirection in place, then *pos += 1. Graph-capturable like the greedy step. */
int32_t janus_whisper_decode_beam_step(const janus_whisper_dims_t* dims, const janus_decoder_weights_t* weights,
                                       const janus_decode_state_t* state, const janus_beam_state_t* beam, void* stream);
/* After the last step: per stream, the pool = finished hypotheses (+ live beams by descending score while it holds
 * fewer than `beam`); best = argmax score / n_tokens. best_tokens int32 [n_streams, max_len] (prompt, sampled tokens,
 * eot padding), best_score [n_streams] sum of log-probs, best_len [n_streams] sampled tokens before eot. */
int32_t janus_whisper_beam_finalize(const janus_decode_state_t* state, const janus_beam_state_t* beam, int32_t eot,
                                    int32_t* best_tokens, float* best_score, int32_t* best_len, void* stream);
/* One layer of decode-time cross-attention for beam search, K/V-cache form: the `beam` rows of a stream (row = stream
 * * beam + k) attend to the same K / V, which one CTA per (stream, head) streams once for all of them.
 * q / out bf16 [n_streams * beam, 64 * n_heads]; 2 <= beam <= JANUS_MAX_BEAM. kv_layout 0: cross_k / cross_v as in
 * janus_dec_cross_attn_bf16 with batch = n_streams (packed-FMA kernel); kv_layout 1: the JANUS_DECODE_KV_SWAPPED
 * layouts, cross_k [n_streams, n_heads, n_ctx, 64] and cross_v [n_streams, n_heads, 64, kv_ld] (tensor-core kernel;
 * kv_ld % 32 == 0; beam >= 1). */
int32_t janus_dec_cross_attn_beam_bf16(const void* q, const void* cross_k, const void* cross_v, void* out,
                                       int32_t n_streams, int32_t beam, int32_t n_heads, int32_t n_ctx, int32_t kv_ld,
                                       int32_t kv_layout, void* stream);
/* janus_whisper_cross_kv_bf16 writing the JANUS_DECODE_KV_SWAPPED layouts. */
int32_t janus_whisper_cross_kv_swapped_bf16(const janus_whisper_dims_t* dims, const janus_decoder_weights_t* weights,
                                            const void* enc_out, int32_t batch, void* cross_k, void* cross_v,
                                            void* stream);
/* janus_dec_xlat_bf16 for rows that share encoder states in groups (beam search): qlat [n_rows, 16, d],
 * enc [n_rows / group, n_ctx, d], z [n_rows, n_heads, d]. */
int32_t janus_dec_xlat_grouped_bf16(const void* qlat, const void* enc, void* z, int32_t n_rows, int32_t group,
                                    int32_t n_ctx, int32_t d, int32_t n_heads, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* JANUS_B200_H */
