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
  int bn_max;  // 0 = auto; otherwise cap the N tile (skinny-M decode GEMMs want many small tiles)
};

int gemm_bf16(const __nv_bfloat16* A, long long lda, long long a_batch_stride, const __nv_bfloat16* W,
              long long ldw, GemmParams p, cudaStream_t stream);
int layernorm_bf16(const __nv_bfloat16* x, const float* w, const float* b, __nv_bfloat16* y, long long rows, int d,
                   long long ldx, long long ldy, cudaStream_t stream);
int attention_enc(const __nv_bfloat16* qkv, __nv_bfloat16* out, int n_batch, int T, int n_heads, cudaStream_t stream);
int num_sms();
int dec_linear(const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, const float* bias,
               const __nv_bfloat16* residual, void* out, int rows, int N, int K, int act, int out_f32,
               cudaStream_t stream);

}  // namespace janus
