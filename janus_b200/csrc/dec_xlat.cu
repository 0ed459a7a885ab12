// Decode-time cross-attention over the ENCODER STATES themselves ("latent" cross-attention).
//
// The reference stack (CTranslate2 Whisper decoder, reached from
// /root/reference/backend/services/transcriber.py:53-57) projects the 1500 encoder states X of a window to
// per-layer keys and values once, K_l = X Wk_l^T and V_l = X Wv_l^T + bv_l, and every decode step reads both:
// 2 * L * 1500 * d bf16 = 55 MB per window and token for Whisper-small -- 55 % of the whole encode step at the
// HBM roofline. Both contractions are linear in X, so they can be moved to the query side exactly:
//
//   scores_h[t] = q_h . K_h[t]      = X[t] . (Wk_h^T q_h)                 =: X[t] . q'_h        (q'_h in R^d)
//   out_h       = sum_t p_h[t] V_h[t] = (sum_t p_h[t] X[t]) Wv_h^T + bv_h =: z_h Wv_h^T + bv_h   (z_h in R^d)
//
// With q' ("absorbed" queries, a grouped GEMM of the 64-wide q_h with Wk_h) and z -> ctx (a grouped GEMM with
// Wv_h) done by the tensor-core GEMM, this kernel only needs X: d * 1500 bf16 = 2.3 MB per window, layer and
// token, HALF the bytes of the K/V formulation (and no per-layer K/V cache at all: 58 GB -> 2.4 GB for 1024
// windows). The price is 12x the attention FLOPs (every head contracts over d = 768 instead of 64), which is
// what tcgen05 is for: per 128-key tile S[128 keys x 16 heads] = X_tile Q'^T (K = d) and
// Y^T[d x 16 heads] += X_tile^T P (K = 128 keys) take ~770 tensor-core cycles against ~7700 cycles of HBM
// time for the tile, so the kernel stays HBM-bound.
//
// One persistent CTA per SM walks windows; 192 threads:
//   warp 0     TMA producer: the window's Q' (d/64 boxes of 16 x 64), then X tile by tile as d/64 boxes of
//              128 keys x 64 features (128B swizzle) into a ring that holds exactly ONE tile (192 KB at d = 768;
//              both GEMMs need the whole tile, so the ring cannot be shorter, and shared memory cannot hold two).
//   warp 1     MMA issuer (convergent, one elected lane): GEMM1 box by box as the boxes land (X box = K-major A,
//              Q' box = K-major B), then, once the softmax warps published P, GEMM2 feature block by feature
//              block (X boxes = MN-major A spanning two swizzle atoms via the descriptor's leading byte offset,
//              P = K-major B), releasing the two boxes of each block to the producer as it completes.
//   warps 2-5  softmax: thread <-> key (TMEM lane). Per tile: S from TMEM, per-head max over the 128 keys
//              (shuffles + one 128-thread named barrier), lazy rescale of Y in TMEM, p = 2^(s - m) to shared
//              memory as the B operand of GEMM2. Per window: Z = Y / l as bf16.
// S is double-buffered in TMEM; TMEM use: 2 x 16 + (d / 128) x 16 columns.
// d <= 768 runs the one-tile-ring kernel described above (dec_xlat_kernel); d = 1024 runs dec_xlat2_kernel, which
// streams every tile twice through a ring of box pairs (second pass from L2) -- see the comment above it.
#include <cstdlib>
#include "kernels.cuh"

namespace janus {

constexpr int XL_KEYS = 128;
constexpr int XL_THREADS = 192;
constexpr int XL_HP = 16;                     // heads padded to the MMA N granularity
constexpr int XL_BOX_BYTES = XL_KEYS * 128;   // 128 keys x 64 bf16 features
constexpr int XL_QBOX_BYTES = XL_HP * 128;    // 16 heads x 64 bf16 features
constexpr int XL_P_BYTES = XL_HP * 256;       // 16 heads x 128 keys bf16 (two 64-key swizzle atoms)
constexpr int XL_MAX_BOXES = 12;              // d <= 768
#ifndef XL_DEFAULT_STREAM
#define XL_DEFAULT_STREAM 0
#endif
constexpr float XL_RESCALE_LOG2 = 8.f;        // lazy rescale threshold (log2 units): p stays below 2^8

static size_t xlat_smem_bytes(int d) {
  const int nb = d / 64;
  return (size_t)nb * (XL_BOX_BYTES + XL_QBOX_BYTES) + 2 * XL_P_BYTES + 1024;
}

__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};\n" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void xl_tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void xl_softmax_barrier() { asm volatile("bar.sync 1, 128;\n" ::: "memory"); }
__device__ __forceinline__ void st_shared_u16(uint32_t addr, uint16_t v) {
  asm volatile("st.shared.u16 [%0], %1;\n" ::"r"(addr), "h"(v) : "memory");
}

#ifdef XL_PROFILE   // phase timing of one CTA (build with JANUS_NVCC_EXTRA=-DXL_PROFILE, scripts/gpu_xlat_bench.py)
__device__ long long xl_dbg[16];
#define XL_T0 long long xl_t = clock64();
#define XL_ACC(acc) { const long long n_ = clock64(); acc += n_ - xl_t; xl_t = n_; }
#else
#define XL_T0
#define XL_ACC(acc)
#endif

// Softmax / epilogue warps (4 warps, thread <-> key == TMEM lane), shared by both kernels below.
__device__ __forceinline__ void xl_softmax_warps(int warp, int lane, uint32_t tmem_base, uint32_t tmem_y, uint64_t* s_full,
                                                 uint64_t* p_full, uint64_t* pv_done, uint64_t* y_full, uint8_t* sP,
                                                 float* red, __nv_bfloat16* z, int T, int d, int n_heads, int n_windows,
                                                 int n_tiles, int nfb, float scale_log2) {
  const int q4 = warp & 3;                 // TMEM lane quarter this warp may access
  const int key = q4 * 32 + lane;          // key within the tile == TMEM lane (== feature row of Y^T blocks)
  const uint32_t lane_off = (uint32_t)(q4 * 32) << 16;
  const int sw = warp - 2;
  float* red_max = red;                    // [2][4][16]
  float* red_sum = red + 128;              // [4][16]
  int g = 0, i = 0;
#ifdef XL_PROFILE
  long long a_ws = 0, a_mx = 0, a_pp = 0, a_ep = 0;
#endif
  XL_T0
  for (int w = blockIdx.x; w < n_windows; w += gridDim.x, ++i) {
    float m_run[XL_HP], l_part[XL_HP];
#pragma unroll
    for (int h = 0; h < XL_HP; ++h) { m_run[h] = -INFINITY; l_part[h] = 0.f; }
    for (int t = 0; t < n_tiles; ++t, ++g) {
      mbar_wait(&s_full[g & 1], (g >> 1) & 1);
      XL_ACC(a_ws)
      tc_fence_after();
      uint32_t sr[XL_HP];
      tmem_ld_32x32b_x16(tmem_base + (g & 1) * XL_HP + lane_off, sr);
      tmem_ld_wait();
      const bool valid = t * XL_KEYS + key < T;
      float s[XL_HP], mx[XL_HP];
#pragma unroll
      for (int h = 0; h < XL_HP; ++h) {
        s[h] = valid ? __uint_as_float(sr[h]) * scale_log2 : -INFINITY;
        // max over the warp's 32 keys with one REDUX per head: floats order like sign-magnitude integers, so
        // flipping the magnitude bits of negatives gives a monotonic int32 key (and back with the same flip)
        int k = __float_as_int(s[h]);
        k ^= (k >> 31) & 0x7fffffff;
        k = __reduce_max_sync(FULL_MASK, k);
        k ^= (k >> 31) & 0x7fffffff;
        mx[h] = __int_as_float(k);
      }
      float* rm = red_max + (g & 1) * 64;
      if (lane < XL_HP) {
        float v = mx[0];
#pragma unroll
        for (int h = 1; h < XL_HP; ++h) v = lane == h ? mx[h] : v;
        rm[sw * XL_HP + lane] = v;
      }
      xl_softmax_barrier();
      XL_ACC(a_mx)
      bool need = t == 0;
#pragma unroll
      for (int h = 0; h < XL_HP; ++h) {
        mx[h] = fmaxf(fmaxf(rm[h], rm[XL_HP + h]), fmaxf(rm[2 * XL_HP + h], rm[3 * XL_HP + h]));
        need |= mx[h] - m_run[h] > XL_RESCALE_LOG2;
      }
      if (need) {   // block-uniform: every thread sees the same maxima
        float alpha[XL_HP];
#pragma unroll
        for (int h = 0; h < XL_HP; ++h) {
          const float m_new = fmaxf(m_run[h], mx[h]);
          alpha[h] = ex2_approx(m_run[h] - m_new);
          m_run[h] = m_new;
          l_part[h] *= alpha[h];
        }
        if (t > 0) {
          mbar_wait(pv_done, (g - 1) & 1);   // GEMM2 of the previous tile has landed in Y
          tc_fence_after();
          for (int fb = 0; fb < nfb; ++fb) {
            uint32_t y[XL_HP];
            tmem_ld_32x32b_x16(tmem_y + fb * XL_HP + lane_off, y);
            tmem_ld_wait();
#pragma unroll
            for (int h = 0; h < XL_HP; ++h) y[h] = __float_as_uint(__uint_as_float(y[h]) * alpha[h]);
            tmem_st_32x32b_x16(tmem_y + fb * XL_HP + lane_off, y);
          }
          xl_tmem_st_wait();
        }
      }
      // p = 2^(s - m) -> bf16 B operand of GEMM2: row = head, 128 keys = two 64-key 128B-swizzled atoms
      const uint32_t pb = smem_u32(sP) + (g & 1) * XL_P_BYTES + (key >> 6) * 2048 + (key & 7) * 2;
      const int kc = (key & 63) >> 3;
#pragma unroll
      for (int h = 0; h < XL_HP; ++h) {
        const float p = ex2_approx(s[h] - m_run[h]);
        l_part[h] += p;
        st_shared_u16(pb + h * 128 + ((kc ^ (h & 7)) << 4), __bfloat16_as_ushort(__float2bfloat16_rn(p)));
      }
      fence_proxy_async_smem();   // generic-proxy smem writes -> visible to the tensor core
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
      XL_ACC(a_pp)
    }
    // ---- window epilogue: Z[w, h, :] = Y[h, :] / l[h]
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int h = 0; h < XL_HP; ++h) l_part[h] += __shfl_xor_sync(FULL_MASK, l_part[h], o);
    }
    if (lane < XL_HP) {
      float v = l_part[0];
#pragma unroll
      for (int h = 1; h < XL_HP; ++h) v = lane == h ? l_part[h] : v;
      red_sum[sw * XL_HP + lane] = v;
    }
    xl_softmax_barrier();
    float inv[XL_HP];
#pragma unroll
    for (int h = 0; h < XL_HP; ++h)
      inv[h] = 1.f / ((red_sum[h] + red_sum[XL_HP + h]) + (red_sum[2 * XL_HP + h] + red_sum[3 * XL_HP + h]));
    mbar_wait(y_full, i & 1);
    tc_fence_after();
    for (int fb = 0; fb < nfb; ++fb) {
      uint32_t y[XL_HP];
      tmem_ld_32x32b_x16(tmem_y + fb * XL_HP + lane_off, y);
      tmem_ld_wait();
      __nv_bfloat16* zp = z + (long long)w * n_heads * d + fb * 128 + key;
#pragma unroll
      for (int h = 0; h < XL_HP; ++h)
        if (h < n_heads) zp[(long long)h * d] = __float2bfloat16_rn(__uint_as_float(y[h]) * inv[h]);
    }
    xl_softmax_barrier();   // red_sum is rewritten by the next window's epilogue only after everyone read it
    XL_ACC(a_ep)
  }
#ifdef XL_PROFILE
  if (blockIdx.x == 3 && threadIdx.x == 64) { xl_dbg[8] = a_ws; xl_dbg[9] = a_mx; xl_dbg[10] = a_pp; xl_dbg[11] = a_ep; xl_dbg[12] = g; }
#endif
}

__global__ void __launch_bounds__(XL_THREADS, 1)
dec_xlat_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_q,
                __nv_bfloat16* z, int T, int d, int n_heads, int n_windows, float scale_log2, int x_group) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const int nb = d >> 6;     // 64-feature boxes per tile
  const int nfb = d >> 7;    // 128-feature blocks of Y^T
  uint8_t* sX = smem;
  uint8_t* sQ = sX + nb * XL_BOX_BYTES;
  uint8_t* sP = sQ + nb * XL_QBOX_BYTES;
  float* red = reinterpret_cast<float*>(sP + 2 * XL_P_BYTES);   // [2][4][16] tile maxima, [4][16] row sums
  uint64_t* bars = reinterpret_cast<uint64_t*>(red + 192);
  uint64_t* box_full = bars;                      // [12]  TMA -> MMA
  uint64_t* box_empty = bars + XL_MAX_BOXES;      // [12]  GEMM2 done with the box -> TMA
  uint64_t* s_full = bars + 2 * XL_MAX_BOXES;     // [2]   GEMM1 -> softmax
  uint64_t* p_full = s_full + 2;                  //       softmax (4 warps) -> GEMM2
  uint64_t* pv_done = p_full + 1;                 //       GEMM2 of the tile complete (Y may be rescaled)
  uint64_t* q_full = pv_done + 1;                 //       Q' of the window landed
  uint64_t* q_empty = q_full + 1;                 //       last GEMM1 of the window complete
  uint64_t* y_full = q_empty + 1;                 //       last GEMM2 of the window complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(y_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_tiles = (T + XL_KEYS - 1) / XL_KEYS;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_q);
    for (int i = 0; i < XL_MAX_BOXES; ++i) {
      mbar_init(&box_full[i], 1);
      mbar_init(&box_empty[i], 1);
    }
    mbar_init(&s_full[0], 1);
    mbar_init(&s_full[1], 1);
    mbar_init(p_full, 4);
    mbar_init(pv_done, 1);
    mbar_init(q_full, 1);
    mbar_init(q_empty, 1);
    mbar_init(y_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_y = tmem_base + 32;   // S buffers at columns [0,16) and [16,32)
  pdl_launch_dependents();
  pdl_wait();   // Q' comes from the previous kernel of the decode chain

  if (warp == 0) {
    // ===================== TMA producer =====================
    const bool leader = elect_one();
    int g = 0;   // tiles issued so far by this CTA (all windows)
    int i = 0;
    for (int w = blockIdx.x; w < n_windows; w += gridDim.x, ++i) {
      if (i > 0) mbar_wait(q_empty, (i - 1) & 1);
      if (leader) {
        mbar_arrive_expect_tx(q_full, nb * XL_QBOX_BYTES);
        for (int kb = 0; kb < nb; ++kb) tma_load_3d(sQ + kb * XL_QBOX_BYTES, &tmap_q, q_full, kb * 64, 0, w);
      }
      __syncwarp();
      for (int t = 0; t < n_tiles; ++t, ++g) {
        for (int kb = 0; kb < nb; ++kb) {
          if (g > 0) mbar_wait(&box_empty[kb], (g - 1) & 1);
          if (leader) {
            mbar_arrive_expect_tx(&box_full[kb], XL_BOX_BYTES);
            tma_load_3d(sX + kb * XL_BOX_BYTES, &tmap_x, &box_full[kb], kb * 64, t * XL_KEYS, w / x_group);
          }
          __syncwarp();
        }
        // The ring holds one tile, so the loads of the next tile cannot be issued before GEMM2 releases the boxes:
        // pull the next tile (and the next window's Q') into L2 now, so that those loads are L2 hits.
        if (leader) {
          int tn = t + 1, wn = w;
          if (tn == n_tiles) { tn = 0; wn = w + gridDim.x; }
          if (wn < n_windows) {
            for (int kb = 0; kb < nb; ++kb) tma_prefetch_3d(&tmap_x, kb * 64, tn * XL_KEYS, wn / x_group);
            if (tn == 0)
              for (int kb = 0; kb < nb; ++kb) tma_prefetch_3d(&tmap_q, kb * 64, 0, wn);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const bool leader = elect_one();
    constexpr uint32_t idesc_s = umma_idesc_bf16(XL_KEYS, XL_HP, 0, 0);
    constexpr uint32_t idesc_y = umma_idesc_bf16(128, XL_HP, 1, 0);   // A = X^T: MN-major
    const uint32_t x_addr = smem_u32(sX), q_addr = smem_u32(sQ), p_addr = smem_u32(sP);
    int g = 0, i = 0;
#ifdef XL_PROFILE
    long long a_box = 0, a_g1 = 0, a_p = 0, a_g2 = 0, a_box0 = 0;
#endif
    XL_T0
    for (int w = blockIdx.x; w < n_windows; w += gridDim.x, ++i) {
      mbar_wait(q_full, i & 1);
      for (int t = 0; t < n_tiles; ++t, ++g) {
        const uint32_t d_s = tmem_base + (g & 1) * XL_HP;
        for (int kb = 0; kb < nb; ++kb) {   // GEMM1: S[128 keys x 16 heads] += X_box[128 x 64] Q'_box[16 x 64]^T
          mbar_wait(&box_full[kb], g & 1);
#ifdef XL_PROFILE
          if (kb == 0) XL_ACC(a_box0) else XL_ACC(a_box)
#endif
          tc_fence_after();
          if (leader) {
            const uint64_t da = umma_smem_desc_sw128(x_addr + kb * XL_BOX_BYTES);
            const uint64_t db = umma_smem_desc_sw128(q_addr + kb * XL_QBOX_BYTES);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_f16(d_s, da + 2 * k, db + 2 * k, idesc_s, (kb | k) != 0);
          }
          __syncwarp();
          XL_ACC(a_g1)
        }
        if (leader) {
          umma_commit(&s_full[g & 1]);
          if (t == n_tiles - 1) umma_commit(q_empty);
        }
        __syncwarp();
        mbar_wait(p_full, g & 1);
        XL_ACC(a_p)
        tc_fence_after();
        for (int fb = 0; fb < nfb; ++fb) {   // GEMM2: Y^T[128 features x 16 heads] += X_boxes^T[128 x 128 keys] P[128 keys x 16]
          if (leader) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {   // 16 keys per MMA = 2048 B of each box
              const uint64_t da = umma_smem_desc_sw128(x_addr + 2 * fb * XL_BOX_BYTES + ks * 2048, 1024, XL_BOX_BYTES);
              const uint64_t db = umma_smem_desc_sw128(p_addr + (g & 1) * XL_P_BYTES + (ks >> 2) * 2048) + 2 * (ks & 3);
              umma_f16(tmem_y + fb * XL_HP, da, db, idesc_y, (t | ks) != 0);
            }
            umma_commit(&box_empty[2 * fb]);
            umma_commit(&box_empty[2 * fb + 1]);
          }
          __syncwarp();
        }
        if (leader) {
          umma_commit(pv_done);
          if (t == n_tiles - 1) umma_commit(y_full);
        }
        __syncwarp();
        XL_ACC(a_g2)
      }
    }
#ifdef XL_PROFILE
    if (blockIdx.x == 3 && lane == 0) { xl_dbg[0] = a_box0; xl_dbg[1] = a_box; xl_dbg[2] = a_g1; xl_dbg[3] = a_p; xl_dbg[4] = a_g2; xl_dbg[5] = g; }
#endif
  } else {
    // ===================== softmax / epilogue (warps 2..5) =====================
    xl_softmax_warps(warp, lane, tmem_base, tmem_y, s_full, p_full, pv_done, y_full, sP, red, z, T, d, n_heads, n_windows,
                     n_tiles, nfb, scale_log2);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

// ------------------------------------------------------------------------------------------------------------
// Streaming variant: the X tile is not held in shared memory. Every tile is read twice, as d/128 box PAIRS
// (2 x 16 KB, adjacent, so that GEMM2's MN-major A operand spans them with LBO = 16 KB) through a ring of pair
// slots: pass A from HBM for GEMM1, pass B again -- an L2 hit, 400 KB per SM later -- for GEMM2. The job
// order is  A(0) A(1) B(0) A(2) B(1) ...  so that GEMM1 of tile g+1 runs on the tensor pipe while the softmax
// warps work on tile g, and loads never wait for the softmax: the one-tile ring of the kernel above serialises
// load -> GEMM1 -> softmax -> GEMM2 -> reload. Q' is double-buffered because A(g+1) may
// already belong to the next window. No limit on d from the ring (Q' and TMEM bound it instead).
// ------------------------------------------------------------------------------------------------------------
constexpr int XL_PAIR_BYTES = 2 * XL_BOX_BYTES;

static int xlat2_pair_slots(int d) {
  const int nb = d / 64;
  const long long fixed = 2LL * nb * XL_QBOX_BYTES + 2 * XL_P_BYTES + 1024;
  int ps = (int)((232448 - fixed) / XL_PAIR_BYTES);
  return ps > 6 ? 6 : ps;
}
static size_t xlat2_smem_bytes(int d) {
  const int nb = d / 64;
  return (size_t)xlat2_pair_slots(d) * XL_PAIR_BYTES + 2 * (size_t)nb * XL_QBOX_BYTES + 2 * XL_P_BYTES + 1024;
}

__global__ void __launch_bounds__(XL_THREADS, 1)
dec_xlat2_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_q,
                 __nv_bfloat16* z, int T, int d, int n_heads, int n_windows, float scale_log2, int PS, int tmem_cols,
                 int x_group) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const int nb = d >> 6;     // 64-feature boxes per tile
  const int nfb = d >> 7;    // box pairs per tile == 128-feature blocks of Y^T
  uint8_t* sX = smem;                                  // PS pair slots
  uint8_t* sQ = sX + PS * XL_PAIR_BYTES;               // 2 x nb boxes
  uint8_t* sP = sQ + 2 * nb * XL_QBOX_BYTES;
  float* red = reinterpret_cast<float*>(sP + 2 * XL_P_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(red + 192);
  uint64_t* full = bars;               // [6]  TMA -> MMA, per pair slot
  uint64_t* empty = bars + 6;          // [6]  MMAs on the slot complete -> TMA
  uint64_t* s_full = bars + 12;        // [2]
  uint64_t* p_full = s_full + 2;
  uint64_t* pv_done = p_full + 1;
  uint64_t* q_full = pv_done + 1;      // [2]
  uint64_t* q_empty = q_full + 2;      // [2]
  uint64_t* y_full = q_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(y_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_tiles = (T + XL_KEYS - 1) / XL_KEYS;
  const int n_win_cta = blockIdx.x < n_windows ? (n_windows - blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int G = n_win_cta * n_tiles;   // tiles this CTA walks

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_q);
    for (int i = 0; i < 6; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
    }
    mbar_init(p_full, 4);
    mbar_init(pv_done, 1);
    mbar_init(y_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_y = tmem_base + 32;   // S buffers at columns [0,16) and [16,32)
  pdl_launch_dependents();
  pdl_wait();   // Q' comes from the previous kernel of the decode chain

  if (warp == 0) {
    // ===================== TMA producer: job order A(0) A(1) B(0) A(2) B(1) ... =====================
    const bool leader = elect_one();
    int job = 0;   // pair jobs issued so far
    for (int step = 0; step <= G; ++step) {
      for (int pass = 0; pass < 2; ++pass) {
        const int g = pass == 0 ? step : step - 1;   // pass 0: A(step), pass 1: B(step - 1)
        if (g < 0 || g >= G) continue;
        const int wi = g / n_tiles, t = g - wi * n_tiles;
        const int w = blockIdx.x + wi * gridDim.x;
        if (pass == 0 && t == 0) {   // first GEMM1 of a window: its Q' into buffer wi & 1
          if (wi >= 2) mbar_wait(&q_empty[wi & 1], ((wi >> 1) - 1) & 1);
          if (leader) {
            mbar_arrive_expect_tx(&q_full[wi & 1], nb * XL_QBOX_BYTES);
            for (int kb = 0; kb < nb; ++kb)
              tma_load_3d(sQ + ((wi & 1) * nb + kb) * XL_QBOX_BYTES, &tmap_q, &q_full[wi & 1], kb * 64, 0, w);
          }
          __syncwarp();
        }
        for (int fb = 0; fb < nfb; ++fb, ++job) {
          const int slot = job % PS;
          if (job >= PS) mbar_wait(&empty[slot], ((job / PS) - 1) & 1);
          if (leader) {
            uint8_t* dst = sX + slot * XL_PAIR_BYTES;
            mbar_arrive_expect_tx(&full[slot], XL_PAIR_BYTES);
            tma_load_3d(dst, &tmap_x, &full[slot], (2 * fb) * 64, t * XL_KEYS, w / x_group);
            tma_load_3d(dst + XL_BOX_BYTES, &tmap_x, &full[slot], (2 * fb + 1) * 64, t * XL_KEYS, w / x_group);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (same job order) =====================
    const bool leader = elect_one();
    constexpr uint32_t idesc_s = umma_idesc_bf16(XL_KEYS, XL_HP, 0, 0);
    constexpr uint32_t idesc_y = umma_idesc_bf16(128, XL_HP, 1, 0);   // A = X^T: MN-major
    const uint32_t x_addr = smem_u32(sX), q_addr = smem_u32(sQ), p_addr = smem_u32(sP);
    int job = 0;
    for (int step = 0; step <= G; ++step) {
      if (step < G) {   // ---- A(step): S[128 keys x 16 heads] = X_tile Q'^T, pair by pair as the boxes land
        const int g = step;
        const int wi = g / n_tiles, t = g - wi * n_tiles;
        if (t == 0) mbar_wait(&q_full[wi & 1], (wi >> 1) & 1);
        const uint32_t d_s = tmem_base + (g & 1) * XL_HP;
        for (int fb = 0; fb < nfb; ++fb, ++job) {
          const int slot = job % PS;
          mbar_wait(&full[slot], (job / PS) & 1);
          tc_fence_after();
          if (leader) {
#pragma unroll
            for (int hb = 0; hb < 2; ++hb) {
              const uint64_t da = umma_smem_desc_sw128(x_addr + slot * XL_PAIR_BYTES + hb * XL_BOX_BYTES);
              const uint64_t db = umma_smem_desc_sw128(q_addr + ((wi & 1) * nb + 2 * fb + hb) * XL_QBOX_BYTES);
#pragma unroll
              for (int k = 0; k < 4; ++k) umma_f16(d_s, da + 2 * k, db + 2 * k, idesc_s, (fb | hb | k) != 0);
            }
            umma_commit(&empty[slot]);
          }
          __syncwarp();
        }
        if (leader) {
          umma_commit(&s_full[g & 1]);
          if (t == n_tiles - 1) umma_commit(&q_empty[wi & 1]);
        }
        __syncwarp();
      }
      if (step >= 1) {   // ---- B(step - 1): Y^T[128 features x 16 heads] += X_pair^T P, once the softmax warps published P
        const int g = step - 1;
        const int t = g % n_tiles;
        mbar_wait(p_full, g & 1);
        tc_fence_after();
        for (int fb = 0; fb < nfb; ++fb, ++job) {
          const int slot = job % PS;
          mbar_wait(&full[slot], (job / PS) & 1);
          tc_fence_after();
          if (leader) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {   // 16 keys per MMA = 2048 B of each box
              const uint64_t da = umma_smem_desc_sw128(x_addr + slot * XL_PAIR_BYTES + ks * 2048, 1024, XL_BOX_BYTES);
              const uint64_t db = umma_smem_desc_sw128(p_addr + (g & 1) * XL_P_BYTES + (ks >> 2) * 2048) + 2 * (ks & 3);
              umma_f16(tmem_y + fb * XL_HP, da, db, idesc_y, (t | ks) != 0);
            }
            umma_commit(&empty[slot]);
          }
          __syncwarp();
        }
        if (leader) {
          umma_commit(pv_done);
          if (t == n_tiles - 1) umma_commit(y_full);
        }
        __syncwarp();
      }
    }
  } else {
    xl_softmax_warps(warp, lane, tmem_base, tmem_y, s_full, p_full, pv_done, y_full, sP, red, z, T, d, n_heads, n_windows,
                     n_tiles, nfb, scale_log2);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

// qlat [n_windows, 16, d] (rows >= n_heads zero), enc [n_windows / x_group, T, d], z [n_windows, n_heads, d].
// x_group > 1 (beam search): x_group consecutive query rows attend to the same encoder states; they are walked by
// neighbouring CTAs at the same time, so that one of them pulls a tile from HBM and the others hit it in L2.
int dec_xlat(const __nv_bfloat16* qlat, const __nv_bfloat16* enc, __nv_bfloat16* z, int n_windows, int T, int d,
             int n_heads, cudaStream_t stream, int x_group) {
  JANUS_CHECK_ARG(x_group >= 1 && n_windows % x_group == 0, "xlat: n_windows=%d is not a multiple of x_group=%d",
                  n_windows, x_group);
  JANUS_CHECK_ARG(d % 128 == 0 && d <= 1024 && n_heads <= XL_HP && n_heads * 64 == d,
                  "xlat: d=%d / n_heads=%d unsupported (need d %% 128 == 0, d <= 1024, <= 16 heads)", d, n_heads);
  CUtensorMap tx, tq;
  {
    const uint64_t dims[3] = {(uint64_t)d, (uint64_t)T, (uint64_t)(n_windows / x_group)};
    const uint64_t str[2] = {(uint64_t)d * 2, (uint64_t)T * d * 2};
    const uint32_t box[3] = {64, XL_KEYS, 1};
    int rc = make_tmap_bf16(&tx, enc, 3, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)d, (uint64_t)XL_HP, (uint64_t)n_windows};
    const uint64_t str[2] = {(uint64_t)d * 2, (uint64_t)XL_HP * d * 2};
    const uint32_t box[3] = {64, XL_HP, 1};
    int rc = make_tmap_bf16(&tq, qlat, 3, dims, str, box);
    if (rc) return rc;
  }
  // JANUS_OPT_XLAT_STREAM: 0 = one-tile ring (d <= 768), 1 = streaming two-pass kernel
  const int variant = option(JANUS_OPT_XLAT_STREAM, XL_DEFAULT_STREAM);
  const bool stream_variant = variant || d / 64 > XL_MAX_BOXES;   // d > 768: a whole tile does not fit the ring
  int max_ctas = num_sms();
  const int cap = option(JANUS_OPT_XLAT_CTAS, 0);   // n > 0 caps the persistent grid (scaling experiments)
  if (cap > 0 && cap < max_ctas) max_ctas = cap;
  const int grid = n_windows < max_ctas ? n_windows : max_ctas;
  const float scale_log2 = 0.125f * 1.4426950408889634f;   // 1/sqrt(64) * log2(e)
  if (stream_variant) {
    const int ps = xlat2_pair_slots(d);
    const int tmem_cols = 32 + (d / 128) * XL_HP <= 128 ? 128 : 256;
    const size_t smem = xlat2_smem_bytes(d);
    static size_t attr2[64] = {};
    const int dev = current_device_slot();
    if (smem > attr2[dev]) {
      JANUS_CUDA(cudaFuncSetAttribute(dec_xlat2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr2[dev] = smem;
    }
    JANUS_CUDA(launch_pdl(dec_xlat2_kernel, dim3(grid), dim3(XL_THREADS), smem, stream, tx, tq, z, T, d, n_heads,
                          n_windows, scale_log2, ps, tmem_cols, x_group));
    return JANUS_OK;
  }
  const size_t smem = xlat_smem_bytes(d);
  static size_t attr_bytes[64] = {};
  const int dev1 = current_device_slot();
  if (smem > attr_bytes[dev1]) {
    JANUS_CUDA(cudaFuncSetAttribute(dec_xlat_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_bytes[dev1] = smem;
  }
  JANUS_CUDA(launch_pdl(dec_xlat_kernel, dim3(grid), dim3(XL_THREADS), smem, stream, tx, tq, z, T, d, n_heads,
                        n_windows, scale_log2, x_group));
  return JANUS_OK;
}

}  // namespace janus

#ifdef XL_PROFILE
extern "C" int32_t janus_debug_xlat_profile(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, janus::xl_dbg, sizeof(long long) * 16) == cudaSuccess ? 0 : -1;
}
#endif

extern "C" int32_t janus_dec_xlat_bf16(const void* qlat, const void* enc, void* z, int32_t n_windows, int32_t n_ctx,
                                       int32_t d, int32_t n_heads, void* stream) {
  JANUS_CHECK_ARG(qlat && enc && z && n_windows > 0 && n_ctx > 0, "xlat: bad argument");
  return janus::dec_xlat(static_cast<const __nv_bfloat16*>(qlat), static_cast<const __nv_bfloat16*>(enc),
                         static_cast<__nv_bfloat16*>(z), n_windows, n_ctx, d, n_heads, (cudaStream_t)stream, 1);
}
extern "C" int32_t janus_dec_xlat_grouped_bf16(const void* qlat, const void* enc, void* z, int32_t n_rows, int32_t group,
                                               int32_t n_ctx, int32_t d, int32_t n_heads, void* stream) {
  JANUS_CHECK_ARG(qlat && enc && z && n_rows > 0 && n_ctx > 0 && group > 0, "xlat: bad argument");
  return janus::dec_xlat(static_cast<const __nv_bfloat16*>(qlat), static_cast<const __nv_bfloat16*>(enc),
                         static_cast<__nv_bfloat16*>(z), n_rows, n_ctx, d, n_heads, (cudaStream_t)stream, group);
}
