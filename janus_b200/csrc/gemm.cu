// K2 building block: bf16 x bf16 -> fp32 GEMM on tcgen05 tensor cores with fused epilogues.
//
//   out[b, t, n] = epilogue( sum_k A[b, t, k] * W[n, k] )       A, W both K-major (row-major, K contiguous)
//
// Used for every dense contraction of the Whisper encoder (CTranslate2 `Whisper.encode` in the
// reference's stack, reached from /root/reference/backend/services/transcriber.py:53-57): the two
// conv1d stem layers (as GEMMs over strided, overlapping TMA views of the time-major activations),
// QKV / output projections, the MLP, and the decoder's per-window cross-K/V projections.
//
// Design (sm_100a): persistent CTAs (one per SM), 320 threads in three roles --
//   warp 0      TMA producer: cp.async.bulk.tensor loads of 128x64 A and BNx64 W tiles (128B swizzle)
//               into a STAGES-deep shared-memory ring, completion via mbarrier complete_tx
//   warp 1      MMA issuer: one thread issues tcgen05.mma (M=128, N=BN, K=16) from shared-memory
//               descriptors into a double-buffered fp32 accumulator in TMEM; tcgen05.commit releases
//               smem stages and publishes finished accumulators
//   warps 2-9   epilogue: tcgen05.ld (32 lanes x 32 columns) -> bias / GELU / positional / residual
//               -> bf16 stores; overlaps the next tile's mainloop through the second TMEM buffer.
#include "kernels.cuh"

namespace janus {

#ifndef GEMM_EPI_PREFETCH
#define GEMM_EPI_PREFETCH 1
#endif
constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_EPI_WARPS = 8;                       // 2 per TMEM lane quarter (column halves)
constexpr int GEMM_THREADS = 64 + 32 * GEMM_EPI_WARPS;

template <int BN>
struct GemmSmem {
  static constexpr int STAGES = BN == 256 ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  static constexpr int B_BYTES = BN * GEMM_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
  // BN >= 128: each epilogue warp owns a 32-row x 128-byte staging slab (128B-swizzled) so that row-major
  // bf16 results / residuals move to and from global memory as full 128-byte lines
  static constexpr int EPI_BYTES = BN >= 128 ? GEMM_EPI_WARPS * 4096 : 0;
  static constexpr int BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

#ifdef GEMM_PROFILE   // phase stamps of CTA 0 (globaltimer ns): build with JANUS_NVCC_EXTRA=-DGEMM_PROFILE
__device__ unsigned long long gemm_dbg[8];
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define GSTAMP(i) if (blockIdx.x == 0 && lane == 0) gemm_dbg[i] = gtime();
#else
#define GSTAMP(i)
#endif

template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const GemmParams p) {
  using S = GemmSmem<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_stage = smem + S::STAGES * S::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_stage + S::EPI_BYTES);
  uint64_t* full = bars;                       // [STAGES]
  uint64_t* empty = bars + S::STAGES;          // [STAGES]
  uint64_t* tmem_full = bars + 2 * S::STAGES;  // [2]
  uint64_t* tmem_empty = tmem_full + 2;        // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (warp == 0) { GSTAMP(0) }

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int i = 0; i < S::STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], GEMM_EPI_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, S::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (warp == 0) { GSTAMP(1) }

  if (warp == 0) {
    // ===================== TMA producer =====================
    // The whole warp walks the loops (convergent code keeps coordinates and descriptors in uniform registers);
    // one elected lane executes the TMA instructions.
    const bool leader = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      const int n_tile = tile % p.n_tiles;
      const int mt = tile / p.n_tiles;
      const int b = mt / p.m_tiles_per_batch;
      const int m_tile = mt % p.m_tiles_per_batch;
      for (int kt = 0; kt < p.k_tiles; ++kt) {
        mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* sa = smem + stage * S::STAGE_BYTES;
        uint8_t* sb = sa + S::A_BYTES;
        if (leader) {
          mbar_arrive_expect_tx(&full[stage], S::STAGE_BYTES);
          tma_load_3d(sa, &tmap_a, &full[stage], kt * GEMM_BK, m_tile * GEMM_BM, b);
          tma_load_2d(sb, &tmap_b, &full[stage], kt * GEMM_BK, n_tile * BN + b * p.w_batch_rows);
        }
        __syncwarp();
        if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // Convergent as well: issuing from inside `if (lane == 0)` made the compiler wrap every tcgen05.mma in a
    // per-thread R2UR / BRA.U.ANY loop.
    const bool leader = elect_one();
    constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int local = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++local) {
      const int as = local & 1;
      const uint32_t aphase = (local >> 1) & 1;
      mbar_wait(&tmem_empty[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      for (int kt = 0; kt < p.k_tiles; ++kt) {
        mbar_wait(&full[stage], phase);
#ifdef GEMM_PROFILE
        if (local == 0 && kt == 0) { GSTAMP(2) }
        if (local == 0 && kt == p.k_tiles - 1) { GSTAMP(3) }
#endif
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * S::STAGE_BYTES);
        const uint64_t da = umma_smem_desc_sw128(sa);
        const uint64_t db = umma_smem_desc_sw128(sa + S::A_BYTES);
        if (leader) {
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k) {
            // +32 bytes (16 bf16) along K inside the 128B swizzle atom = +2 in the >>4 address field
            umma_f16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kt | k) != 0);
          }
          umma_commit(&empty[stage]);
          if (kt == p.k_tiles - 1) umma_commit(&tmem_full[as]);
        }
        __syncwarp();
        if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    const int q = warp & 3;            // TMEM lane quarter this warp may access
    const int chalf = (warp - 2) >> 2; // which half of the BN columns this warp drains
    int local = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++local) {
      const int n_tile = tile % p.n_tiles;
      const int mt = tile / p.n_tiles;
      const int b = mt / p.m_tiles_per_batch;
      const int m_tile = mt % p.m_tiles_per_batch;
      const int as = local & 1;
      const uint32_t aphase = (local >> 1) & 1;
      const int t = m_tile * GEMM_BM + q * 32 + lane;  // row within the batch
      const bool row_ok = t < p.M_per_batch;
      const long long row_off = (long long)b * p.out_batch_stride + (long long)(t + p.out_row_offset) * p.ldc;
      const float* bias = p.bias ? p.bias + (long long)b * p.bias_batch_stride : nullptr;   // grouped GEMM: per-batch bias
      // BN == 64 is the skinny-M (decode) configuration: one tile per CTA, so nothing hides the epilogue's
      // dependent global loads (measured 3.4-4.5 us of a 10-11 us kernel). Request bias and residual while
      // the mainloop is still running.
      float4 bb_pre[8];
      uint4 res_pre[4];
      if constexpr (GEMM_EPI_PREFETCH && BN == 64) {
        const int n0p = n_tile * BN + chalf * 32;
        if (bias) {
#pragma unroll
          for (int i = 0; i < 8; ++i) bb_pre[i] = __ldg(reinterpret_cast<const float4*>(bias + n0p) + i);
        }
        if (p.residual && row_ok) {
#pragma unroll
          for (int i = 0; i < 4; ++i) res_pre[i] = __ldg(reinterpret_cast<const uint4*>(p.residual + row_off + n0p) + i);
        }
      }
      mbar_wait(&tmem_full[as], aphase);
#ifdef GEMM_PROFILE
      if (local == 0 && warp == 2) { GSTAMP(4) }
#endif
      tc_fence_after();
      const uint32_t taddr = tmem_base + as * BN + ((uint32_t)(q * 32) << 16);
      if (BN >= 128 && p.store_mode == STORE_ROWMAJOR) {
        // ---- coalesced path: 64 output columns (one 128-byte line per row) at a time through the warp's
        // swizzled staging slab. A thread owns row `lane` while computing (TMEM lane = row) and the
        // 16-byte chunk (lane & 7) of rows 4*it + (lane >> 3) while talking to global memory.
        uint8_t* slab = epi_stage + (warp - 2) * 4096;
        const uint32_t slab_u = smem_u32(slab);
        const int grow0 = m_tile * GEMM_BM + q * 32;                   // first row of this warp within the batch
        const long long base_off = (long long)b * p.out_batch_stride + (long long)(grow0 + p.out_row_offset) * p.ldc;
        const int cr = lane >> 3, cc = lane & 7;
#pragma unroll 1
        for (int g0 = chalf * (BN / 2); g0 < (chalf + 1) * (BN / 2); g0 += 64) {
          const int n0 = n_tile * BN + g0;
          if (p.residual) {   // global (full lines) -> slab
            uint4 ru[8];
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const int rr = 4 * it + cr;
              ru[it] = make_uint4(0, 0, 0, 0);
              if (grow0 + rr < p.M_per_batch)
                ru[it] = __ldg(reinterpret_cast<const uint4*>(p.residual + base_off + (long long)rr * p.ldc + n0) + cc);
            }
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const int rr = 4 * it + cr;
              *reinterpret_cast<uint4*>(slab + rr * 128 + ((cc ^ (rr & 7)) << 4)) = ru[it];
            }
            __syncwarp();
          }
#pragma unroll
          for (int hc = 0; hc < 2; ++hc) {   // two 32-column halves of the 64-column group
            uint32_t r[32];
            tmem_ld_32x32b_x32(taddr + g0 + 32 * hc, r);
            float4 bb[8];
            if (bias) {
#pragma unroll
              for (int i = 0; i < 8; ++i) bb[i] = __ldg(reinterpret_cast<const float4*>(bias + n0 + 32 * hc) + i);
            }
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
            if (bias) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                v[4 * i] += bb[i].x; v[4 * i + 1] += bb[i].y; v[4 * i + 2] += bb[i].z; v[4 * i + 3] += bb[i].w;
              }
            }
            if (p.act == ACT_GELU) {
#pragma unroll
              for (int i = 0; i < 32; i += 2) gelu_erf_fast2(v[i], v[i + 1]);
            }
            if (p.pos && row_ok) {
              const float* pp = p.pos + (long long)t * p.N + n0 + 32 * hc;
#pragma unroll
              for (int i = 0; i < 32; i += 4) {
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(pp + i));
                v[i] += b4.x; v[i + 1] += b4.y; v[i + 2] += b4.z; v[i + 3] += b4.w;
              }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {   // this row's chunks 4*hc + i
              const uint32_t addr = slab_u + lane * 128 + (((4 * hc + i) ^ (lane & 7)) << 4);
              if (p.residual) {
                uint4 u;
                asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "r"(addr) : "memory");
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const float2 f = __bfloat1622float2(h[j]);
                  v[i * 8 + 2 * j] += f.x;
                  v[i * 8 + 2 * j + 1] += f.y;
                }
              }
              uint4 o;
              __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
              for (int j = 0; j < 4; ++j) ho[j] = __floats2bfloat162_rn(v[i * 8 + 2 * j], v[i * 8 + 2 * j + 1]);
              asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
            }
          }
          __syncwarp();
#pragma unroll
          for (int it = 0; it < 8; ++it) {   // slab -> global, 4 full 128-byte lines per instruction
            const int rr = 4 * it + cr;
            const uint4 u = *reinterpret_cast<const uint4*>(slab + rr * 128 + ((cc ^ (rr & 7)) << 4));
            if (grow0 + rr < p.M_per_batch)
              *(reinterpret_cast<uint4*>(p.out + base_off + (long long)rr * p.ldc + n0) + cc) = u;
          }
          __syncwarp();
        }
      } else
#pragma unroll 1
      for (int c0 = chalf * (BN / 2); c0 < (chalf + 1) * (BN / 2); c0 += 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr + c0, r);
        tmem_ld_wait();
        const int n0 = n_tile * BN + c0;
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
        if (bias) {
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            float4 bb;
            if constexpr (GEMM_EPI_PREFETCH && BN == 64) bb = bb_pre[i >> 2];
            else bb = __ldg(reinterpret_cast<const float4*>(bias + n0 + i));
            v[i] += bb.x; v[i + 1] += bb.y; v[i + 2] += bb.z; v[i + 3] += bb.w;
          }
        }
        if (p.act == ACT_GELU) {
#pragma unroll
          for (int i = 0; i < 32; i += 2) gelu_erf_fast2(v[i], v[i + 1]);
        }
        if (row_ok) {
          if (p.pos) {
            const float* pp = p.pos + (long long)t * p.N + n0;
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float4 bb = __ldg(reinterpret_cast<const float4*>(pp + i));
              v[i] += bb.x; v[i + 1] += bb.y; v[i + 2] += bb.z; v[i + 3] += bb.w;
            }
          }
          if (p.residual) {
            const uint4* rp = reinterpret_cast<const uint4*>(p.residual + row_off + n0);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 u;
              if constexpr (GEMM_EPI_PREFETCH && BN == 64) u = res_pre[i];
              else u = __ldg(rp + i);
              const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float2 f = __bfloat1622float2(h[j]);
                v[i * 8 + 2 * j] += f.x;
                v[i * 8 + 2 * j + 1] += f.y;
              }
            }
          }
          if (p.store_mode == STORE_F32) {
            float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + row_off + n0);
#pragma unroll
            for (int i = 0; i < 8; ++i) op[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          } else if (p.store_mode == STORE_ROWMAJOR) {
            uint4* op = reinterpret_cast<uint4*>(p.out + row_off + n0);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 u;
              __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
              for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(v[i * 8 + 2 * j], v[i * 8 + 2 * j + 1]);
              op[i] = u;
            }
          } else {
            // cross-attention K/V cache layouts (see decoder.cu)
            const int l = n0 / (2 * p.kv_d);
            const int rem = n0 - l * 2 * p.kv_d;
            const int is_v = rem >= p.kv_d;
            const int f0 = rem - is_v * p.kv_d;
            const int h = f0 >> 6, dim0 = f0 & 63;
            const long long bh = ((long long)l * p.n_batch + b) * p.kv_heads + h;
            __nv_bfloat16* base = is_v ? p.kv_v : p.kv_k;
            if (is_v != p.kv_swap) {   // row-major [t][64] (V by default, K with kv_swap)
              uint4* op = reinterpret_cast<uint4*>(base + (bh * p.M_per_batch + t) * 64 + dim0);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                uint4 u;
                __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
                for (int j = 0; j < 4; ++j) hh[j] = __floats2bfloat162_rn(v[i * 8 + 2 * j], v[i * 8 + 2 * j + 1]);
                op[i] = u;
              }
            } else {                   // transposed [64][kv_ld], positions contiguous (K by default, V with kv_swap)
              __nv_bfloat16* op = base + (bh * 64 + dim0) * p.kv_ld + t;
#pragma unroll
              for (int i = 0; i < 32; ++i) op[(long long)i * p.kv_ld] = __float2bfloat16_rn(v[i]);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[as]);
    }
  }

  if (warp == 2) { GSTAMP(5) }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, S::TMEM_COLS);
    GSTAMP(6)
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(ptr);
  }
  return fn;
}

int make_tmap_bf16(CUtensorMap* out, const void* gptr, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled driver entry point unavailable");
    return JANUS_ERR_CUDA;
  }
  cuuint64_t gdim[5];
  cuuint64_t gstr[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
  }
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(gptr), gdim, gstr,
                   bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d dims %llu,%llu,%llu)", (int)r, rank,
              (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
              (unsigned long long)(rank > 2 ? dims[2] : 0));
    return JANUS_ERR_CUDA;
  }
  return JANUS_OK;
}

// per-device caches: kernel attributes and the SM count belong to the CURRENT device (a process may drive several)
int current_device_slot() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev >= 0 && dev < 64 ? dev : 63;
}
int num_sms() {
  static int n_sms[64] = {};
  const int dev = current_device_slot();
  if (!n_sms[dev]) {
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    n_sms[dev] = v > 0 ? v : 148;
  }
  return n_sms[dev];
}

template <int BN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, GemmParams& p, cudaStream_t stream) {
  using S = GemmSmem<BN>;
  p.n_tiles = p.N / BN;
  p.total_tiles = p.n_batch * p.m_tiles_per_batch * p.n_tiles;
  static bool attr_set[64] = {};
  const int dev = current_device_slot();
  if (!attr_set[dev]) {
    JANUS_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::BYTES));
    attr_set[dev] = true;
  }
  const int grid = p.total_tiles < num_sms() ? p.total_tiles : num_sms();
  gemm_bf16_kernel<BN><<<grid, GEMM_THREADS, S::BYTES, stream>>>(ta, tb, p);
  JANUS_LAUNCH_CHECK("gemm_bf16_kernel");
  return JANUS_OK;
}

// Internal entry used by the encoder/decoder drivers and by the public ABI below.
int gemm_bf16(const __nv_bfloat16* A, long long lda, long long a_batch_stride, const __nv_bfloat16* W,
              long long ldw, GemmParams p, cudaStream_t stream) {
  JANUS_CHECK_ARG(p.N % 64 == 0, "gemm: N=%d must be a multiple of 64", p.N);
  JANUS_CHECK_ARG(p.K % 8 == 0 && lda % 8 == 0 && ldw % 8 == 0 && a_batch_stride % 8 == 0,
                  "gemm: K and leading dimensions must be multiples of 8 elements (16 B)");
  JANUS_CHECK_ARG((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0,
                  "gemm: operands must be 16 B aligned");
  p.m_tiles_per_batch = (p.M_per_batch + GEMM_BM - 1) / GEMM_BM;
  p.k_tiles = (p.K + GEMM_BK - 1) / GEMM_BK;
  int BN = (p.N % 256 == 0) ? 256 : (p.N % 128 == 0 ? 128 : 64);
  if (p.bn_max > 0 && BN > p.bn_max) BN = p.bn_max >= 128 ? 128 : 64;
  CUtensorMap ta, tb;
  {
    const uint64_t dims[3] = {(uint64_t)p.K, (uint64_t)p.M_per_batch, (uint64_t)p.n_batch};
    const uint64_t str[2] = {(uint64_t)lda * 2, (uint64_t)(p.n_batch > 1 ? a_batch_stride : (long long)p.M_per_batch * lda) * 2};
    const uint32_t box[3] = {GEMM_BK, GEMM_BM, 1};
    int rc = make_tmap_bf16(&ta, A, 3, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)p.K, (uint64_t)p.N + (uint64_t)(p.n_batch - 1) * p.w_batch_rows};
    const uint64_t str[1] = {(uint64_t)ldw * 2};
    const uint32_t box[2] = {GEMM_BK, (uint32_t)BN};
    int rc = make_tmap_bf16(&tb, W, 2, dims, str, box);
    if (rc) return rc;
  }
  if (BN == 256) return launch_gemm<256>(ta, tb, p, stream);
  if (BN == 128) return launch_gemm<128>(ta, tb, p, stream);
  return launch_gemm<64>(ta, tb, p, stream);
}

}  // namespace janus

#ifdef GEMM_PROFILE
extern "C" int32_t janus_debug_gemm_profile(unsigned long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, janus::gemm_dbg, sizeof(unsigned long long) * 8) == cudaSuccess ? 0 : -1;
}
#endif

using namespace janus;

extern "C" int32_t janus_gemm_bf16(const void* A, int64_t lda, int64_t a_batch_stride, const void* W, int64_t ldw,
                                   void* out, int64_t ldc, int64_t out_batch_stride, int32_t out_row_offset,
                                   int32_t M_per_batch, int32_t n_batch, int32_t N, int32_t K, const float* bias,
                                   const void* residual, const float* pos, int32_t act, void* stream) {
  JANUS_CHECK_ARG(A && W && out, "gemm: null pointer");
  JANUS_CHECK_ARG(M_per_batch > 0 && n_batch > 0 && N > 0 && K > 0, "gemm: bad shape");
  JANUS_CHECK_ARG(ldc % 8 == 0 && out_batch_stride % 8 == 0, "gemm: output strides must be multiples of 8");
  GemmParams p = {};
  p.M_per_batch = M_per_batch; p.n_batch = n_batch; p.N = N; p.K = K;
  p.out = static_cast<__nv_bfloat16*>(out); p.ldc = ldc; p.out_batch_stride = out_batch_stride;
  p.out_row_offset = out_row_offset; p.bias = bias;
  p.residual = static_cast<const __nv_bfloat16*>(residual); p.pos = pos; p.act = act;
  p.store_mode = STORE_ROWMAJOR;
  return gemm_bf16(static_cast<const __nv_bfloat16*>(A), lda, a_batch_stride, static_cast<const __nv_bfloat16*>(W),
                   ldw, p, (cudaStream_t)stream);
}
