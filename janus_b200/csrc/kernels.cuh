// Internal (non-ABI) launch helpers shared between the kernel translation units and the model drivers.
#pragma once
#include "tc_common.cuh"

namespace janus {

enum GemmAct { ACT_NONE = 0, ACT_GELU = 1 };
enum GemmStore { STORE_ROWMAJOR = 0, STORE_CROSS_KV = 1, STORE_F32 = 2 };

struct GemmParams {
  int M_per_batch, n_batch, N, K;
  int m_tiles_per_batch, n_tiles, total_tiles, k_tiles;
  __nv_bfloat16* out;
  long long ldc, out_batch_stride;
  int out_row_offset;
  const float* bias;                  // [N] or null
  const __nv_bfloat16* residual;      // indexed like out, or null
  const float* pos;                   // [M_per_batch, N] added after the activation, or null
  int act;
  int store_mode;
  // STORE_CROSS_KV: N = n_layer * 2 * d; columns [l][k|v][h][64]. K goes to kT[l][b][h][64][kv_ld]
  // (keys contiguous), V to v[l][b][h][t][64].
  __nv_bfloat16* kv_k;
  __nv_bfloat16* kv_v;
  int kv_d, kv_heads, kv_ld;
  int kv_swap;  // 1: the two layouts swap roles -- K row-major k[l][b][h][t][64], V transposed v[l][b][h][64][kv_ld]
  int bn_max;  // 0 = auto; otherwise cap the N tile (skinny-M decode GEMMs want many small tiles)
  // grouped GEMM: batch b multiplies by W rows [b * w_batch_rows, b * w_batch_rows + N) and adds
  // bias[b * bias_batch_stride + n]; 0 = every batch shares W / bias
  int w_batch_rows, bias_batch_stride;
};

int gemm_bf16(const __nv_bfloat16* A, long long lda, long long a_batch_stride, const __nv_bfloat16* W,
              long long ldw, GemmParams p, cudaStream_t stream);
int layernorm_bf16(const __nv_bfloat16* x, const float* w, const float* b, __nv_bfloat16* y, long long rows, int d,
                   long long ldx, long long ldy, cudaStream_t stream, bool pdl = false);
int attention_enc(const __nv_bfloat16* qkv, __nv_bfloat16* out, int n_batch, int T, int n_heads, cudaStream_t stream);
int num_sms();
// latent cross-attention (dec_xlat.cu): qlat [n, 16, d] absorbed queries, enc [n, T, d] encoder states -> z [n, H, d]
// (x_group > 1: x_group consecutive query rows share one window of encoder states -- beam search)
int dec_xlat(const __nv_bfloat16* qlat, const __nv_bfloat16* enc, __nv_bfloat16* z, int n_windows, int T, int d,
             int n_heads, cudaStream_t stream, int x_group = 1);
int dec_linear(const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, const float* bias,
               const __nv_bfloat16* residual, void* out, int rows, int N, int K, int act, int out_f32,
               cudaStream_t stream, bool pdl = false);

// ---- beam search bookkeeping (dec_beam.cu) ------------------------------------------------------
struct BeamStepParams {
  const float* logits;        // [rows, n_vocab_pad]
  const uint8_t* suppress;
  int32_t* tokens;            // [rows, max_len]
  int32_t* pos;
  int32_t* n_done;
  int n_vocab, n_vocab_pad, max_len, n_prompt, eot, blank, ts_begin, max_initial_ts, self_ld;
  janus_beam_state_t bm;
};
// candidates (rule filter + log-softmax + top beam+1 per row), then per stream: selection, finished list, in-place
// reorder of token histories / indirection
int dec_beam_step(const BeamStepParams& p, cudaStream_t stream);
int dec_beam_finalize(const int32_t* tokens, const int32_t* pos, int max_len, int n_prompt, int eot,
                      const janus_beam_state_t& bm, int32_t* best_tokens, float* best_score, int32_t* best_len,
                      cudaStream_t stream);


}  // namespace janus
