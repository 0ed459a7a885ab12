// K2: fused (flash-style) non-causal multi-head attention for the Whisper encoder on tcgen05.
//
//   out[b, t, h*64 : h*64+64] = softmax( q_h k_h^T / sqrt(64) ) v_h        (head_dim = 64 for every Whisper size)
//
// Input is the packed QKV activation [B, T, 3d] written by the QKV GEMM; Whisper scales q and k by
// d_h^-1/4 each (CTranslate2 `Whisper.encode` MultiHeadAttention), which is the 1/sqrt(d_h) applied to
// the scores here.
//
// One CTA = 256 queries of one (batch, head), as two 128-row tiles owned by two softmax warpgroups
// that ping-pong on the tensor core (FlashAttention-4 style schedule):
//   warp 8 (one thread)  TMA: Q tiles once, K/V tiles through a 3-stage ring (128B swizzle)
//   warp 9 (one thread)  tcgen05.mma issue:  S_w = Q_w K_j^T  (M128 N128 K16 x4 -> TMEM cols w*128)
//                                             O_w += P_w V_j   (M128 N64  K16 x8 -> TMEM cols 256+w*64;
//                        V is consumed straight from its row-major TMA tile as an MN-major B operand).
//                        S_w(j+1) is issued as soon as warpgroup w has published P_w(j), so one
//                        warpgroup's softmax always overlaps the other's MMAs.
//   warps 0-3 / 4-7      one thread per query row (TMEM lane): tcgen05.ld of the 128 scores, thread-local
//                        row max / row sum (no shuffles), P -> bf16 -> shared memory in the K-major
//                        128B-swizzled UMMA layout. O stays in TMEM and accumulates across key tiles; it is
//                        rescaled (tcgen05.ld/st) only when the running max moved by more than 2^8
//                        ("lazy rescale"), which removes the O round trip from almost every tile.
#include "kernels.cuh"

namespace janus {

constexpr int ATT_BQ = 256;   // queries per CTA: two 128-row tiles, one per softmax warpgroup
constexpr int ATT_BK = 128;
constexpr int ATT_DH = 64;
constexpr int ATT_KV_STAGES = 3;
constexpr int ATT_THREADS = 320;  // warps 0-3 softmax WG0, 4-7 softmax WG1, 8 TMA, 9 MMA
constexpr int ATT_TILE_BYTES = 128 * 64 * 2;  // 16 KB: a Q, K or V tile
// Q0 Q1 | (K V) x stages | P0 P1 (32 KB each) | barriers
constexpr int ATT_SMEM = ATT_TILE_BYTES * (2 + 2 * ATT_KV_STAGES + 4) + 256;
constexpr float ATT_RESCALE_LOG2 = 8.0f;  // lazy rescale threshold: p = exp2(.) stays below 2^8

__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};\n" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory");
}

__global__ void __launch_bounds__(ATT_THREADS, 1)
attention_enc_kernel(const __grid_constant__ CUtensorMap tmap_qkv, __nv_bfloat16* __restrict__ out, int T, int d,
                     float scale_log2) {
  extern __shared__ __align__(1024) uint8_t smem[];  // 128B-swizzled tiles need 1024 B alignment
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* sQ[2] = {smem, smem + ATT_TILE_BYTES};
  uint8_t* sKV = smem + 2 * ATT_TILE_BYTES;                       // stage s: K at +2s, V at +2s+1 tiles
  uint8_t* sP[2] = {sKV + 2 * ATT_KV_STAGES * ATT_TILE_BYTES, sKV + (2 * ATT_KV_STAGES + 2) * ATT_TILE_BYTES};
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + (2 * ATT_KV_STAGES + 4) * ATT_TILE_BYTES);
  uint64_t* q_full = bars;                    // [1]
  uint64_t* kv_full = bars + 1;               // [stages]
  uint64_t* kv_empty = kv_full + ATT_KV_STAGES;
  uint64_t* s_full = kv_empty + ATT_KV_STAGES;  // [2]
  uint64_t* p_full = s_full + 2;                // [2]
  uint64_t* o_ready = p_full + 2;               // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_ready + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * ATT_BQ;
  const int h = blockIdx.y;
  const int b = blockIdx.z;
  const int n_tiles = (T + ATT_BK - 1) / ATT_BK;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_qkv);
    mbar_init(q_full, 1);
    for (int i = 0; i < ATT_KV_STAGES; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
      mbar_init(&o_ready[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 9) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // TMEM columns: S0 [0,128)  S1 [128,256)  O0 [256,320)  O1 [320,384)

  if (warp == 8) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, 2 * ATT_TILE_BYTES);
      tma_load_3d(sQ[0], &tmap_qkv, q_full, h * ATT_DH, q0, b);
      tma_load_3d(sQ[1], &tmap_qkv, q_full, h * ATT_DH, q0 + 128, b);
      for (int j = 0; j < n_tiles; ++j) {
        const int s = j % ATT_KV_STAGES;
        const uint32_t ph = (j / ATT_KV_STAGES) & 1;
        mbar_wait(&kv_empty[s], ph ^ 1);
        uint8_t* k = sKV + 2 * s * ATT_TILE_BYTES;
        mbar_arrive_expect_tx(&kv_full[s], 2 * ATT_TILE_BYTES);
        tma_load_3d(k, &tmap_qkv, &kv_full[s], d + h * ATT_DH, j * ATT_BK, b);
        tma_load_3d(k + ATT_TILE_BYTES, &tmap_qkv, &kv_full[s], 2 * d + h * ATT_DH, j * ATT_BK, b);
      }
    }
  } else if (warp == 9) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, 0, 1);  // B (= V) is MN-major
      const uint64_t dq[2] = {umma_smem_desc_sw128(smem_u32(sQ[0])), umma_smem_desc_sw128(smem_u32(sQ[1]))};
      const uint32_t p_addr[2] = {smem_u32(sP[0]), smem_u32(sP[1])};
      auto issue_s = [&](int w, int j) {   // S_w = Q_w K_j^T
        const uint64_t dk = umma_smem_desc_sw128(smem_u32(sKV + 2 * (j % ATT_KV_STAGES) * ATT_TILE_BYTES));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16(tmem_base + w * 128, dq[w] + 2 * k, dk + 2 * k, idesc_s, k != 0);
        umma_commit(&s_full[w]);
      };
      auto issue_pv = [&](int w, int j) {  // O_w (+)= P_w V_j
        const uint32_t v_addr = smem_u32(sKV + (2 * (j % ATT_KV_STAGES) + 1) * ATT_TILE_BYTES);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t dp = umma_smem_desc_sw128(p_addr[w] + (ks >> 2) * ATT_TILE_BYTES) + 2 * (ks & 3);
          const uint64_t dv = umma_smem_desc_sw128(v_addr + ks * 2048);
          umma_f16(tmem_base + 256 + w * 64, dp, dv, idesc_pv, (j | ks) != 0);
        }
        umma_commit(&o_ready[w]);
      };
      mbar_wait(q_full, 0);
      mbar_wait(&kv_full[0], 0);
      tc_fence_after();
      issue_s(0, 0);
      issue_s(1, 0);
      for (int j = 0; j < n_tiles; ++j) {
        const bool more = j + 1 < n_tiles;
        if (more) {
          mbar_wait(&kv_full[(j + 1) % ATT_KV_STAGES], ((j + 1) / ATT_KV_STAGES) & 1);
        }
#pragma unroll
        for (int w = 0; w < 2; ++w) {
          mbar_wait(&p_full[w], j & 1);
          tc_fence_after();
          issue_pv(w, j);
          if (w == 1) umma_commit(&kv_empty[j % ATT_KV_STAGES]);  // both PV(j) issued: K_j/V_j free when they retire
          if (more) issue_s(w, j + 1);   // S_w is free: softmax w has consumed it before signalling P
        }
      }
    }
  } else {
    // ===================== softmax warpgroups: one thread per query row =====================
    const int w = warp >> 2;                 // warpgroup 0 / 1
    const int r = threadIdx.x & 127;         // row within the 128-query tile == TMEM lane
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tmem_s = tmem_base + w * 128 + lane_off;
    const uint32_t tmem_o = tmem_base + 256 + w * 64 + lane_off;
    float m_used = -INFINITY, l = 0.f;
    uint8_t* p_row = sP[w] + (r >> 3) * 1024 + (r & 7) * 128;
    const int rx = r & 7;
    for (int j = 0; j < n_tiles; ++j) {
      const int n_valid = min(ATT_BK, T - j * ATT_BK);
      mbar_wait(&s_full[w], j & 1);
      tc_fence_after();
      uint32_t s[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld_32x32b_x32(tmem_s + 32 * c, s[c]);
      tmem_ld_wait();
      float mx = -INFINITY;
      if (n_valid == ATT_BK) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(s[c][i]));
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (32 * c + i < n_valid) mx = fmaxf(mx, __uint_as_float(s[c][i]));
      }
      // lazy rescale (warp-uniform, because the TMEM round trip is warp-collective)
      bool need = (j == 0) || ((mx - m_used) * scale_log2 > ATT_RESCALE_LOG2);
      need = __any_sync(FULL_MASK, need);
      if (need) {
        const float m_new = fmaxf(m_used, mx);
        if (j > 0) {
          const float alpha = exp2f((m_used - m_new) * scale_log2);
          l *= alpha;
          mbar_wait(&o_ready[w], (j - 1) & 1);   // PV(j-1) has landed in O
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            uint32_t o[32];
            tmem_ld_32x32b_x32(tmem_o + 32 * c, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32b_x32(tmem_o + 32 * c, o);
          }
          tmem_st_wait();
        }
        m_used = m_new;
      }
      const float mb = m_used * scale_log2;
      float rowsum = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t packed[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = exp2f(fmaf(__uint_as_float(s[c][i]), scale_log2, -mb));
          float p1 = exp2f(fmaf(__uint_as_float(s[c][i + 1]), scale_log2, -mb));
          if (n_valid != ATT_BK) {
            if (32 * c + i >= n_valid) p0 = 0.f;
            if (32 * c + i + 1 >= n_valid) p1 = 0.f;
          }
          rowsum += p0 + p1;
          __nv_bfloat162 hh = __floats2bfloat162_rn(p0, p1);
          packed[i >> 1] = *reinterpret_cast<uint32_t*>(&hh);
        }
        uint8_t* blk = p_row + (c >> 1) * ATT_TILE_BYTES;   // 64-key block of the K-major P tile
        const int cbase = (c & 1) * 4;                       // first 16-byte chunk of this 32-key group
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          uint4 u = make_uint4(packed[cc * 4], packed[cc * 4 + 1], packed[cc * 4 + 2], packed[cc * 4 + 3]);
          *reinterpret_cast<uint4*>(blk + (((cbase + cc) ^ rx) << 4)) = u;
        }
      }
      l += rowsum;
      fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[w]);
    }
    // ---- epilogue: O / l
    mbar_wait(&o_ready[w], (n_tiles - 1) & 1);
    tc_fence_after();
    const int t = q0 + w * 128 + r;
    const float inv = 1.f / l;
    __nv_bfloat16* orow = out + ((long long)b * T + t) * d + h * ATT_DH;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t o[32];
      tmem_ld_32x32b_x32(tmem_o + 32 * c, o);
      tmem_ld_wait();
      if (t < T) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 u;
          __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
          for (int jx = 0; jx < 4; ++jx)
            hh[jx] = __floats2bfloat162_rn(__uint_as_float(o[i * 8 + 2 * jx]) * inv,
                                           __uint_as_float(o[i * 8 + 2 * jx + 1]) * inv);
          reinterpret_cast<uint4*>(orow + 32 * c)[i] = u;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
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
