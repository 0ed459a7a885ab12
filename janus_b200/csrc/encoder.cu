// Whisper encoder driver: sequences the sm_100a kernels for one batch of log-mel windows.
//
// Replaces CTranslate2 `Whisper.encode(features)` as called by faster-whisper's
// WhisperModel.transcribe, which is what /root/reference/backend/services/transcriber.py:53-57 runs.
//   conv1(k3,p1)+GELU and conv2(k3,s2,p1)+GELU(+positional) as tcgen05 GEMMs over overlapping TMA row
//   views of the time-major activations, then n_layers x [LN -> QKV GEMM -> fused attention ->
//   out-proj GEMM (+residual) -> LN -> fc1 GEMM (+GELU) -> fc2 GEMM (+residual)], then ln_post.
// Activations are bf16, accumulation fp32, the residual stream is bf16.
#include "kernels.cuh"

using namespace janus;

static int64_t enc_ws_elems(const janus_whisper_dims_t* c, int64_t B) {
  const int64_t BT = B * c->n_ctx;
  // x, xn, qkv(3), ff(4) -- the stem's h1 [B, 2T+2, d] aliases ff
  const int64_t ff = BT * 4 * c->d;
  const int64_t h1 = B * (2 * (int64_t)c->n_ctx + 2) * c->d;
  return BT * c->d * 5 + (ff > h1 ? ff : h1);
}

extern "C" int64_t janus_whisper_encode_workspace_bytes(const janus_whisper_dims_t* dims, int32_t batch) {
  return enc_ws_elems(dims, batch) * 2 + 256;
}

extern "C" int32_t janus_whisper_encode_bf16(const janus_whisper_dims_t* c, const janus_encoder_weights_t* w,
                                             const void* mel_pad, int32_t batch, void* enc_out, void* workspace,
                                             int64_t workspace_bytes, void* stream_) {
  JANUS_CHECK_ARG(c && w && mel_pad && enc_out && workspace, "encode: null pointer");
  JANUS_CHECK_ARG(batch > 0, "encode: batch must be positive");
  JANUS_CHECK_ARG(c->d == c->n_heads * 64, "encode: head_dim must be 64 (d=%d heads=%d)", c->d, c->n_heads);
  JANUS_CHECK_ARG(c->n_mels % 8 == 0 && c->d % 64 == 0, "encode: n_mels %% 8 and d %% 64 required");
  if (workspace_bytes < janus_whisper_encode_workspace_bytes(c, batch)) {
    set_error("encode: workspace too small (%lld < %lld)", (long long)workspace_bytes,
              (long long)janus_whisper_encode_workspace_bytes(c, batch));
    return JANUS_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream_;
  const int B = batch, T = c->n_ctx, d = c->d, C = c->n_mels;
  const long long BT = (long long)B * T;
  const int Tin = 2 * T;
  __nv_bfloat16* base = reinterpret_cast<__nv_bfloat16*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  __nv_bfloat16* x = base;
  __nv_bfloat16* xn = x + BT * d;
  __nv_bfloat16* qkv = xn + BT * d;
  __nv_bfloat16* ff = qkv + BT * 3 * d;
  __nv_bfloat16* h1 = ff;
  const __nv_bfloat16* mel = static_cast<const __nv_bfloat16*>(mel_pad);
  int rc;

  // conv padding rows of h1 (row 0 and row Tin+1 of every batch) must be zero
  const size_t pitch = (size_t)(Tin + 2) * d * 2;
  JANUS_CUDA(cudaMemset2DAsync(h1, pitch, 0, (size_t)d * 2, B, st));
  JANUS_CUDA(cudaMemset2DAsync(h1 + (long long)(Tin + 1) * d, pitch, 0, (size_t)d * 2, B, st));

  GemmParams p = {};
  // conv1: rows t of the padded mel [Tin+2, C] viewed as overlapping [Tin, 3C]
  p.M_per_batch = Tin; p.n_batch = B; p.N = d; p.K = 3 * C;
  p.out = h1; p.ldc = d; p.out_batch_stride = (long long)(Tin + 2) * d; p.out_row_offset = 1;
  p.bias = w->conv1_b; p.act = ACT_GELU; p.store_mode = STORE_ROWMAJOR;
  if ((rc = gemm_bf16(mel, C, (long long)(Tin + 2) * C, static_cast<const __nv_bfloat16*>(w->conv1_w), 3 * C, p, st))) return rc;
  // conv2 (stride 2): rows 2t..2t+2 of h1 viewed as [T, 3d] with row stride 2d; + positional embedding
  p = GemmParams{};
  p.M_per_batch = T; p.n_batch = B; p.N = d; p.K = 3 * d;
  p.out = x; p.ldc = d; p.out_batch_stride = (long long)T * d; p.out_row_offset = 0;
  p.bias = w->conv2_b; p.act = ACT_GELU; p.pos = w->pos; p.store_mode = STORE_ROWMAJOR;
  if ((rc = gemm_bf16(h1, 2 * d, (long long)(Tin + 2) * d, static_cast<const __nv_bfloat16*>(w->conv2_w), 3 * d, p, st))) return rc;

  auto flat = [&](__nv_bfloat16* out, int N, int K, const float* bias, const __nv_bfloat16* res, int act) {
    GemmParams q = {};
    q.M_per_batch = (int)BT; q.n_batch = 1; q.N = N; q.K = K;
    q.out = out; q.ldc = N; q.out_batch_stride = BT * N; q.bias = bias; q.residual = res; q.act = act;
    q.store_mode = STORE_ROWMAJOR;
    return q;
  };
  for (int l = 0; l < c->n_layers; ++l) {
    const janus_enc_layer_t& L = w->layers[l];
    if ((rc = layernorm_bf16(x, L.ln1_w, L.ln1_b, xn, BT, d, d, d, st))) return rc;
    if ((rc = gemm_bf16(xn, d, BT * d, static_cast<const __nv_bfloat16*>(L.w_qkv), d, flat(qkv, 3 * d, d, L.b_qkv, nullptr, ACT_NONE), st))) return rc;
    if ((rc = attention_enc(qkv, xn, B, T, c->n_heads, st))) return rc;
    if ((rc = gemm_bf16(xn, d, BT * d, static_cast<const __nv_bfloat16*>(L.w_o), d, flat(x, d, d, L.b_o, x, ACT_NONE), st))) return rc;
    if ((rc = layernorm_bf16(x, L.ln2_w, L.ln2_b, xn, BT, d, d, d, st))) return rc;
    if ((rc = gemm_bf16(xn, d, BT * d, static_cast<const __nv_bfloat16*>(L.w_fc1), d, flat(ff, 4 * d, d, L.b_fc1, nullptr, ACT_GELU), st))) return rc;
    if ((rc = gemm_bf16(ff, 4 * d, BT * 4 * d, static_cast<const __nv_bfloat16*>(L.w_fc2), 4 * d, flat(x, d, 4 * d, L.b_fc2, x, ACT_NONE), st))) return rc;
  }
  return layernorm_bf16(x, w->ln_post_w, w->ln_post_b, static_cast<__nv_bfloat16*>(enc_out), BT, d, d, d, st);
}
