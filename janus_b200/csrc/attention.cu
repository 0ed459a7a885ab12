// K2: fused (flash-style) non-causal multi-head attention for the Whisper encoder on tcgen05.
//
//   out[b, t, h*64 : h*64+64] = softmax( q_h k_h^T / sqrt(64) ) v_h        (head_dim = 64 for every Whisper size)
//
// Input is the packed QKV activation [B, T, 3d] written by the QKV GEMM; Whisper scales q and k by
// d_h^-1/4 each (CTranslate2 `Whisper.encode` MultiHeadAttention), which is the 1/sqrt(d_h) applied to
// the scores here.
//
// One CTA = 128 queries of one (batch, head). Per 128-key tile:
//   warp 4 (one thread)  TMA loads K/V tiles (double buffered, 128B swizzle) and issues
//                        S = Q K^T        tcgen05.mma M128 N128 K16 x4   -> TMEM columns [0,128)
//                        PV = P V         tcgen05.mma M128 N64  K16 x8   -> TMEM columns [128,192)
//                        (V is consumed straight from its row-major TMA tile as an MN-major B operand)
//   warps 0-3            one thread per query row (TMEM lane): tcgen05.ld S, online softmax with
//                        thread-local row max / row sum (no shuffles), P -> bf16 -> shared memory in
//                        the K-major 128B-swizzled UMMA layout, then O = O * alpha + PV in registers.
// Two CTAs are resident per SM (112 KB smem, 256 TMEM columns each) so one CTA's softmax overlaps
// the other's MMAs.
#include "kernels.cuh"

namespace janus {

constexpr int ATT_BQ = 128;
constexpr int ATT_BK = 128;
constexpr int ATT_DH = 64;
constexpr int ATT_THREADS = 160;
constexpr int ATT_TILE_BYTES = 128 * 64 * 2;  // 16 KB: Q, K or V tile
constexpr int ATT_SMEM = ATT_TILE_BYTES * 5 + 2 * ATT_TILE_BYTES /*P*/ + 128;  // 2 CTAs/SM must fit in 228 KB

__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_enc_kernel(const __grid_constant__ CUtensorMap tmap_qkv, __nv_bfloat16* __restrict__ out, int T, int d,
                     float scale_log2) {
  extern __shared__ __align__(1024) uint8_t smem[];  // 128B-swizzled tiles need 1024 B alignment
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* sQ = smem;
  uint8_t* sK[2] = {smem + ATT_TILE_BYTES, smem + 3 * ATT_TILE_BYTES};
  uint8_t* sV[2] = {smem + 2 * ATT_TILE_BYTES, smem + 4 * ATT_TILE_BYTES};
  uint8_t* sP = smem + 5 * ATT_TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 7 * ATT_TILE_BYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;   // [2]
  uint64_t* kv_empty = bars + 3;  // [2]
  uint64_t* s_full = bars + 5;
  uint64_t* p_full = bars + 6;
  uint64_t* pv_full = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * ATT_BQ;
  const int h = blockIdx.y;
  const int b = blockIdx.z;
  const int n_tiles = (T + ATT_BK - 1) / ATT_BK;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_qkv);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(p_full, 4);
    mbar_init(pv_full, 1);
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_s = tmem_base;
  const uint32_t tmem_pv = tmem_base + 128;

  if (warp == 4) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, 0, 1);  // B (= V) is MN-major
      mbar_arrive_expect_tx(q_full, ATT_TILE_BYTES);
      tma_load_3d(sQ, &tmap_qkv, q_full, h * ATT_DH, q0, b);
      for (int j = 0; j < 2 && j < n_tiles; ++j) {
        mbar_arrive_expect_tx(&kv_full[j], 2 * ATT_TILE_BYTES);
        tma_load_3d(sK[j], &tmap_qkv, &kv_full[j], d + h * ATT_DH, j * ATT_BK, b);
        tma_load_3d(sV[j], &tmap_qkv, &kv_full[j], 2 * d + h * ATT_DH, j * ATT_BK, b);
      }
      mbar_wait(q_full, 0);
      const uint64_t dq = umma_smem_desc_sw128(smem_u32(sQ));
      const uint32_t p_addr = smem_u32(sP);
      for (int j = 0; j < n_tiles; ++j) {
        const int buf = j & 1;
        mbar_wait(&kv_full[buf], (j >> 1) & 1);
        tc_fence_after();
        const uint64_t dk = umma_smem_desc_sw128(smem_u32(sK[buf]));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16(tmem_s, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
        umma_commit(s_full);
        if (j >= 1 && j + 1 < n_tiles) {
          // refill the other buffer (last read by PV of tile j-1) with tile j+1
          const int ob = buf ^ 1;
          mbar_wait(&kv_empty[ob], ((j - 1) >> 1) & 1);
          mbar_arrive_expect_tx(&kv_full[ob], 2 * ATT_TILE_BYTES);
          tma_load_3d(sK[ob], &tmap_qkv, &kv_full[ob], d + h * ATT_DH, (j + 1) * ATT_BK, b);
          tma_load_3d(sV[ob], &tmap_qkv, &kv_full[ob], 2 * d + h * ATT_DH, (j + 1) * ATT_BK, b);
        }
        mbar_wait(p_full, j & 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV[buf]);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t dp = umma_smem_desc_sw128(p_addr + (ks >> 2) * ATT_TILE_BYTES) + 2 * (ks & 3);
          const uint64_t dv = umma_smem_desc_sw128(v_addr + ks * 2048);
          umma_f16(tmem_pv, dp, dv, idesc_pv, ks != 0);
        }
        umma_commit(pv_full);
        umma_commit(&kv_empty[buf]);
      }
    }
  } else {
    // ===================== softmax / output: one thread per query row =====================
    const int r = threadIdx.x;  // 0..127 == TMEM lane
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    float m = -INFINITY, l = 0.f;
    float O[ATT_DH];
#pragma unroll
    for (int i = 0; i < ATT_DH; ++i) O[i] = 0.f;
    uint8_t* p_row = sP + (r >> 3) * 1024 + (r & 7) * 128;
    const int rx = r & 7;
    for (int j = 0; j < n_tiles; ++j) {
      const int n_valid = min(ATT_BK, T - j * ATT_BK);
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      // pass 1: row max over the valid keys
      float mx = -INFINITY;
#pragma unroll 1
      for (int c0 = 0; c0 < ATT_BK; c0 += 32) {
        if (c0 >= n_valid) break;
        uint32_t s[32];
        tmem_ld_32x32b_x32(tmem_s + lane_off + c0, s);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (c0 + i < n_valid) mx = fmaxf(mx, __uint_as_float(s[i]));
      }
      const float m_new = fmaxf(m, mx);
      const float alpha = exp2f((m - m_new) * scale_log2);  // m = -inf on the first tile -> 0
      const float mb = m_new * scale_log2;
      float rowsum = 0.f;
      // pass 2: p = exp2(s*c - m*c) -> bf16 -> swizzled K-major smem
#pragma unroll 1
      for (int c0 = 0; c0 < ATT_BK; c0 += 32) {
        uint32_t s[32];
        if (c0 < n_valid) {
          tmem_ld_32x32b_x32(tmem_s + lane_off + c0, s);
          tmem_ld_wait();
        }
        uint32_t packed[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = 0.f, p1 = 0.f;
          if (c0 + i < n_valid) p0 = exp2f(fmaf(__uint_as_float(s[i]), scale_log2, -mb));
          if (c0 + i + 1 < n_valid) p1 = exp2f(fmaf(__uint_as_float(s[i + 1]), scale_log2, -mb));
          rowsum += p0 + p1;
          __nv_bfloat162 hh = __floats2bfloat162_rn(p0, p1);
          packed[i >> 1] = *reinterpret_cast<uint32_t*>(&hh);
        }
        uint8_t* blk = p_row + (c0 >> 6) * ATT_TILE_BYTES;
        const int cbase = (c0 & 63) >> 3;  // first 16-byte chunk of this 32-key group within the 64-key block
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          uint4 u = make_uint4(packed[cc * 4], packed[cc * 4 + 1], packed[cc * 4 + 2], packed[cc * 4 + 3]);
          *reinterpret_cast<uint4*>(blk + (((cbase + cc) ^ rx) << 4)) = u;
        }
      }
      l = l * alpha + rowsum;
      m = m_new;
      fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
#pragma unroll
      for (int i = 0; i < ATT_DH; ++i) O[i] *= alpha;
      mbar_wait(pv_full, j & 1);
      tc_fence_after();
#pragma unroll
      for (int c0 = 0; c0 < ATT_DH; c0 += 32) {
        uint32_t pv[32];
        tmem_ld_32x32b_x32(tmem_pv + lane_off + c0, pv);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) O[c0 + i] += __uint_as_float(pv[i]);
      }
      tc_fence_before();
    }
    const int t = q0 + r;
    if (t < T) {
      const float inv = 1.f / l;
      uint4* op = reinterpret_cast<uint4*>(out + ((long long)b * T + t) * d + h * ATT_DH);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        uint4 u;
        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
        for (int jx = 0; jx < 4; ++jx)
          hh[jx] = __floats2bfloat162_rn(O[i * 8 + 2 * jx] * inv, O[i * 8 + 2 * jx + 1] * inv);
        op[i] = u;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

int attention_enc(const __nv_bfloat16* qkv, __nv_bfloat16* out, int n_batch, int T, int n_heads, cudaStream_t stream) {
  const int d = n_heads * ATT_DH;
  CUtensorMap tm;
  const uint64_t dims[3] = {(uint64_t)3 * d, (uint64_t)T, (uint64_t)n_batch};
  const uint64_t str[2] = {(uint64_t)3 * d * 2, (uint64_t)T * 3 * d * 2};
  const uint32_t box[3] = {ATT_DH, 128, 1};
  int rc = make_tmap_bf16(&tm, qkv, 3, dims, str, box);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    JANUS_CUDA(cudaFuncSetAttribute(attention_enc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM));
    attr_set = true;
  }
  const dim3 grid((T + ATT_BQ - 1) / ATT_BQ, n_heads, n_batch);
  const float scale_log2 = 0.125f * 1.4426950408889634f;  // 1/sqrt(64) * log2(e)
  attention_enc_kernel<<<grid, ATT_THREADS, ATT_SMEM, stream>>>(tm, out, T, d, scale_log2);
  JANUS_LAUNCH_CHECK("attention_enc_kernel");
  return JANUS_OK;
}

}  // namespace janus

using namespace janus;

extern "C" int32_t janus_attention_enc_bf16(const void* qkv, void* out, int32_t n_batch, int32_t T, int32_t n_heads,
                                            void* stream) {
  JANUS_CHECK_ARG(qkv && out, "attention: null pointer");
  JANUS_CHECK_ARG(n_batch > 0 && T > 0 && n_heads > 0, "attention: bad shape");
  return attention_enc(static_cast<const __nv_bfloat16*>(qkv), static_cast<__nv_bfloat16*>(out), n_batch, T, n_heads,
                       (cudaStream_t)stream);
}
