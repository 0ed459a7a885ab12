// K3, beam search bookkeeping (BASELINE configs[3]: beam 5).
//
// Replaces the beam search CTranslate2 runs inside `Whisper.generate(beam_size=5, patience=1, length_penalty=1)`
// when faster-whisper is called with beam_size=5 (the reference's own call site,
// /root/reference/backend/services/transcriber.py:53-57, passes beam_size=1 = decoder.cu's greedy pick). The
// algorithm is the published one of OpenAI Whisper's BeamSearchDecoder / MaximumLikelihoodRanker, stated step by
// step in oracle/whisper_ref.py (beam_decode, beam_step_select):
//
//   per live beam (row): rule-filtered log-softmax, candidates = its top beam+1 tokens, score = cum + log-prob;
//   per stream: walk the candidates by descending score (ties: beam, then token ascending; -inf = no candidate):
//     eot candidates finish their hypothesis (kept while the stream holds < max_candidates), the first `beam`
//     others become the next beams; a stream is done when max_candidates hypotheses are finished;
//   at the end: pool = finished (+ live beams by descending score while < beam), best = argmax score / n_tokens.
//
// Top beam+1 per row then merge per stream is exactly the walk over the flat beam * vocab list: a beam has one eot
// candidate, so the walk never goes deeper than beam+1 into any single beam.
//
// Layout decisions: a row's token history is rewritten in place on a reorder (a few hundred bytes per row), but the
// self-attention K/V caches (2 * L * t * d bf16 per row) are NEVER copied: indir[row][j] names the slot of the
// stream that holds position j of the row's hypothesis and dec_self_attn_kernel<true> gathers through it.
#include "kernels.cuh"

namespace janus {

constexpr int BC_THREADS = 512;
constexpr int BC_KC = JANUS_MAX_BEAM + 1;   // candidates per row, compile-time bound

struct RuleState {   // SuppressBlank / ApplyTimestampRules of one row (the rules of dec_pick_kernel) as two index ranges:
  int t_lo, t_hi;    // allowed text tokens   [t_lo, t_hi)  within [0, timestamp_begin)
  int s_lo, s_hi;    // allowed timestamps    [s_lo, s_hi)  within [timestamp_begin, n_vocab)
};

__device__ __forceinline__ RuleState rule_state(const int32_t* toks, int nxt, int n_prompt, int tb, int max_initial_ts,
                                                int eot, int n_vocab) {
  const int n_gen = nxt - n_prompt;
  const int last = n_gen >= 1 ? toks[nxt - 1] : -1;
  const int penult = n_gen >= 2 ? toks[nxt - 2] : -1;
  const bool last_ts = n_gen >= 1 && last >= tb;
  const bool penult_ts = n_gen < 2 || penult >= tb;
  int ts_floor = tb, last_stamp = -1;   // timestamps never decrease
  for (int i = n_prompt; i < nxt; ++i) {
    const int t = toks[i];
    if (t >= tb) last_stamp = t;
  }
  if (last_stamp >= 0) ts_floor = (last_ts && !penult_ts) ? last_stamp : last_stamp + 1;
  RuleState r;
  r.t_lo = 0; r.t_hi = tb; r.s_lo = ts_floor; r.s_hi = n_vocab;
  if (n_gen == 0) {            // first sampled token: a timestamp no later than max_initial_timestamp (blank, eot < tb)
    r.t_hi = 0;
    r.s_lo = tb;
    r.s_hi = tb + max_initial_ts + 1;
  }
  if (last_ts) {               // timestamps come in pairs
    if (penult_ts) r.s_hi = r.s_lo;   // ... <ts><ts>: text only
    else r.t_lo = eot;                // ... <text><ts>: a timestamp or eot-and-above, no normal text
  }
  if (r.s_hi < r.s_lo) r.s_hi = r.s_lo;
  if (r.t_hi < r.t_lo) r.t_hi = r.t_lo;
  return r;
}

__device__ __forceinline__ bool rule_allows(const RuleState& r, int i, bool suppressed) {
  return !suppressed && ((unsigned)(i - r.t_lo) < (unsigned)(r.t_hi - r.t_lo) || (unsigned)(i - r.s_lo) < (unsigned)(r.s_hi - r.s_lo));
}

struct MaxSum {   // running max m and sum 2^(y - m) of y = logit * log2(e); m starts at a large negative FINITE value so
  float m, s;     // that masked tokens (y = -inf) and the first token need no branch (2^(-inf) = 0, no inf - inf)
};
constexpr float MS_EMPTY = -1e30f;
__device__ __forceinline__ float ms_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void ms_add4(MaxSum& a, float y0, float y1, float y2, float y3) {
  const float m4 = fmaxf(fmaxf(y0, y1), fmaxf(y2, y3));
  if (m4 > a.m) {   // rare after the first few tokens
    a.s *= ms_ex2(a.m - m4);
    a.m = m4;
  }
  a.s += (ms_ex2(y0 - a.m) + ms_ex2(y1 - a.m)) + (ms_ex2(y2 - a.m) + ms_ex2(y3 - a.m));
}
__device__ __forceinline__ void ms_merge(MaxSum& a, float om, float os) {
  if (os == 0.f) return;
  if (a.s == 0.f) { a.m = om; a.s = os; return; }
  if (om > a.m) {
    a.s = a.s * ms_ex2(a.m - om) + os;
    a.m = om;
  } else {
    a.s += os * ms_ex2(om - a.m);
  }
}

__device__ __forceinline__ uint32_t load_suppress4(const uint8_t* sup, int i0, int n_vocab) {
  if (i0 + 3 < n_vocab) return *reinterpret_cast<const uint32_t*>(sup + i0);
  uint32_t s4 = 0x01010101u;
  for (int e = 0; e < 4 && i0 + e < n_vocab; ++e) s4 = (s4 & ~(0xffu << (8 * e))) | ((uint32_t)sup[i0 + e] << (8 * e));
  return s4;
}

// (value, index) order of the candidate lists: larger value first, then smaller index
__device__ __forceinline__ bool cand_before(float v, int i, float bv, int bi) { return v > bv || (v == bv && i < bi); }

// ------------------------------------------------------------------------------------------
// candidates: one CTA per row, ONE pass over the 209 KB of fp32 logits. Every thread keeps max / sum-exp of its
// allowed text and timestamp tokens; the top-kc lists are kept PER WARP in shared memory, guarded by a warp-uniform
// threshold (the list's current kc-th entry): a token enters the insertion path only if it beats the threshold,
// which after the first few hundred tokens almost never happens (~kc ln(n / kc) insertions per list), so the scan
// stays convergent. (Per-thread register lists made every warp run the 9-slot insertion code on nearly every
// token -- some lane always had an insertion -- and cost 0.9-1.1 ms per step at 2560 rows.) Two lists per warp --
// all allowed tokens / timestamps only -- because the "timestamps outweigh every text token" rule, which decides
// whether text is masked and which normaliser applies, is only known after the block reduction.
// ------------------------------------------------------------------------------------------
constexpr int BC_WARPS = BC_THREADS / 32;

struct WarpList {   // sorted by (value desc, index asc)
  float v[BC_KC];
  int i[BC_KC];
};

// lanes with `pred` insert (x, idx) into the warp's list one after the other; returns with thr = the kc-th entry
__device__ __forceinline__ void warp_list_offer(WarpList& wl, int kc, bool pred, float x, int idx, float& thr_v,
                                                int& thr_i, int lane) {
  unsigned pending = __ballot_sync(FULL_MASK, pred);
  while (pending) {
    const int src = __ffs(pending) - 1;
    if (lane == src) {
      int j = kc - 1;   // shift entries that come after (x, idx) down by one
      while (j > 0 && cand_before(x, idx, wl.v[j - 1], wl.i[j - 1])) {
        wl.v[j] = wl.v[j - 1];
        wl.i[j] = wl.i[j - 1];
        --j;
      }
      wl.v[j] = x;
      wl.i[j] = idx;
    }
    __syncwarp();
    thr_v = wl.v[kc - 1];
    thr_i = wl.i[kc - 1];
    pending &= pending - 1;
    // a lane whose token no longer beats the raised threshold drops out
    pending &= __ballot_sync(FULL_MASK, pred && cand_before(x, idx, thr_v, thr_i));
    __syncwarp();
  }
}

// order-preserving float <-> int map (for atomicMax on the block-wide thresholds)
__device__ __forceinline__ int f2ord(float f) {
  int k = __float_as_int(f);
  return k ^ ((k >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float ord2f(int k) { return __int_as_float(k ^ ((k >> 31) & 0x7fffffff)); }

__global__ void __launch_bounds__(BC_THREADS, 3) dec_beam_cand_kernel(const BeamStepParams p) {
  __shared__ float sh_m[2][BC_WARPS], sh_s[2][BC_WARPS];
  // block-wide value thresholds: the largest kc-th entry any warp holds. kc tokens >= it exist in the row, so a token
  // below it can never be among the row's top kc -- with it a warp enters the insertion path ~once per row instead
  // of once per iteration (a warp alone sees only 1/16 of the row and its own threshold rises slowly)
  __shared__ int thr_blk[2];
  __shared__ int sh_wmax[BC_WARPS];
  __shared__ WarpList lists[2][BC_WARPS];   // [all allowed | timestamps only][warp]
  __shared__ RuleState sh_rs;
  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  pdl_launch_dependents();
  pdl_wait();
  const int nxt = *p.pos + 1;
  if (nxt >= p.max_len || nxt < p.n_prompt) return;
  const int K = p.bm.beam, kc = K + 1;
  if (p.bm.stream_done[row / K]) return;
  if (tid == 0)
    sh_rs = rule_state(p.tokens + (long long)row * p.max_len, nxt, p.n_prompt, p.ts_begin, p.max_initial_ts, p.eot, p.n_vocab);
  if (tid < 2) thr_blk[tid] = f2ord(-INFINITY);
  if (lane < BC_KC) {
    lists[0][wid].v[lane] = lists[1][wid].v[lane] = -INFINITY;
    lists[0][wid].i[lane] = lists[1][wid].i[lane] = 0x7fffffff;
  }
  __syncthreads();
  const RuleState rs = sh_rs;
  const float* lg = p.logits + (long long)row * p.n_vocab_pad;
  const int tb = p.ts_begin;
  {
    // Warm start of the block-wide threshold: the kc-th largest of the 16 warps' maxima over the row's first 4096
    // tokens. Those maxima are 16 different allowed tokens, so kc tokens >= it exist. Without it every warp fills
    // its list by ~50 serial insertions (60 % of the kernel's instructions); with it ~5.
    int best = f2ord(-INFINITY);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int i0 = tid * 4 + h * BC_THREADS * 4;
      if (i0 < p.n_vocab) {
        const float4 x4 = *reinterpret_cast<const float4*>(lg + i0);
        const uint32_t s4 = load_suppress4(p.suppress, i0, p.n_vocab);
        const float xs[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (i0 + e < p.n_vocab && rule_allows(rs, i0 + e, ((s4 >> (8 * e)) & 0xffu) != 0)) best = max(best, f2ord(xs[e]));
      }
    }
    best = __reduce_max_sync(FULL_MASK, best);
    if (lane == 0) sh_wmax[wid] = best;
    __syncthreads();
    if (tid == 0) {
      int v[BC_WARPS];
      for (int w = 0; w < BC_WARPS; ++w) v[w] = sh_wmax[w];
      for (int r = 0; r < kc; ++r) {   // selection of the kc largest (kc <= 9 <= 16 warps)
        int b = r;
        for (int w = r + 1; w < BC_WARPS; ++w) if (v[w] > v[b]) b = w;
        const int t_ = v[r]; v[r] = v[b]; v[b] = t_;
      }
      thr_blk[0] = v[kc - 1];   // -inf if fewer than kc warps saw an allowed token
    }
    __syncthreads();
  }
  MaxSum text = {MS_EMPTY, 0.f}, ts = {MS_EMPTY, 0.f};
  float thr_av = -INFINITY, thr_tv = -INFINITY;
  int thr_ai = 0x7fffffff, thr_ti = 0x7fffffff;
  WarpList& la = lists[0][wid];
  WarpList& lt = lists[1][wid];
  constexpr int STEP = BC_THREADS * 4;
  constexpr float LOG2E = 1.4426950408889634f;
  for (int i0 = tid * 4; i0 - tid * 4 < p.n_vocab; i0 += 2 * STEP) {   // warp-uniform trip count (votes inside)
    // n_vocab_pad is a multiple of 256 and rows are zero-padded: a load that starts below n_vocab never leaves the row
    const int i1 = i0 + STEP;
    const bool has0 = i0 < p.n_vocab, has1 = i1 < p.n_vocab;
    const float4 xa = has0 ? *reinterpret_cast<const float4*>(lg + i0) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 xb = has1 ? *reinterpret_cast<const float4*>(lg + i1) : make_float4(0.f, 0.f, 0.f, 0.f);
    const uint32_t sa = has0 ? load_suppress4(p.suppress, i0, p.n_vocab) : 0x01010101u;
    const uint32_t sb = has1 ? load_suppress4(p.suppress, i1, p.n_vocab) : 0x01010101u;
    const float xs[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
    // fast path (branch-free per token): range filter, statistics per group of 4, and ONE vote per list on whether
    // any of the 8 tokens reaches the threshold max(block-wide, this warp's); the insertion path runs only then
    const float gate_a = fmaxf(ord2f(*reinterpret_cast<volatile int*>(&thr_blk[0])), thr_av);
    const float gate_t = fmaxf(ord2f(*reinterpret_cast<volatile int*>(&thr_blk[1])), thr_tv);
    unsigned okm = 0;
    float y[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int i = (e < 4 ? i0 : i1) + (e & 3);
      const uint32_t sup = ((e < 4 ? sa : sb) >> (8 * (e & 3))) & 0xffu;
      const bool ok = i < p.n_vocab && rule_allows(rs, i, sup != 0);
      okm |= (unsigned)ok << e;
      y[e] = ok ? xs[e] * LOG2E : -INFINITY;
    }
    unsigned hit_a = 0, hit_t = 0;
#pragma unroll
    for (int g4 = 0; g4 < 2; ++g4) {
      const int ib = g4 ? i1 : i0;
      const float* yy = y + 4 * g4;
      unsigned tsm;   // which of the group's 4 tokens are timestamps
      if (ib + 3 < tb) {
        ms_add4(text, yy[0], yy[1], yy[2], yy[3]);
        tsm = 0;
      } else if (ib >= tb) {
        ms_add4(ts, yy[0], yy[1], yy[2], yy[3]);
        tsm = 0xf;
      } else {   // the one group of the row that straddles timestamp_begin
        tsm = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) tsm |= (unsigned)(ib + e >= tb) << e;
        ms_add4(text, (tsm & 1) ? -INFINITY : yy[0], (tsm & 2) ? -INFINITY : yy[1], (tsm & 4) ? -INFINITY : yy[2],
                (tsm & 8) ? -INFINITY : yy[3]);
        ms_add4(ts, (tsm & 1) ? yy[0] : -INFINITY, (tsm & 2) ? yy[1] : -INFINITY, (tsm & 4) ? yy[2] : -INFINITY,
                (tsm & 8) ? yy[3] : -INFINITY);
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const unsigned ok = (okm >> (4 * g4 + e)) & 1u;
        const float x = xs[4 * g4 + e];
        hit_a |= (ok & (unsigned)(x >= gate_a)) << (4 * g4 + e);
        hit_t |= (ok & ((tsm >> e) & 1u) & (unsigned)(x >= gate_t)) << (4 * g4 + e);
      }
    }
    if (__any_sync(FULL_MASK, hit_a != 0)) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int i = (e < 4 ? i0 : i1) + (e & 3);
        warp_list_offer(la, kc, ((hit_a >> e) & 1u) && cand_before(xs[e], i, thr_av, thr_ai), xs[e], i, thr_av, thr_ai, lane);
      }
      if (lane == 0) atomicMax(&thr_blk[0], f2ord(thr_av));
    }
    if (__any_sync(FULL_MASK, hit_t != 0)) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int i = (e < 4 ? i0 : i1) + (e & 3);
        warp_list_offer(lt, kc, ((hit_t >> e) & 1u) && cand_before(xs[e], i, thr_tv, thr_ti), xs[e], i, thr_tv, thr_ti, lane);
      }
      if (lane == 0) atomicMax(&thr_blk[1], f2ord(thr_tv));
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ms_merge(text, __shfl_xor_sync(FULL_MASK, text.m, o), __shfl_xor_sync(FULL_MASK, text.s, o));
    ms_merge(ts, __shfl_xor_sync(FULL_MASK, ts.m, o), __shfl_xor_sync(FULL_MASK, ts.s, o));
  }
  if (lane == 0) { sh_m[0][wid] = text.m; sh_s[0][wid] = text.s; sh_m[1][wid] = ts.m; sh_s[1][wid] = ts.s; }
  __syncthreads();
  text.m = sh_m[0][0]; text.s = sh_s[0][0]; ts.m = sh_m[1][0]; ts.s = sh_s[1][0];
  for (int w = 1; w < BC_WARPS; ++w) { ms_merge(text, sh_m[0][w], sh_s[0][w]); ms_merge(ts, sh_m[1][w], sh_s[1][w]); }
  // back to natural-log units: max text logit, log-sum-exp of the timestamps / of everything allowed
  const float text_max = text.s == 0.f ? -INFINITY : text.m / LOG2E;
  const float lse_ts = ts.s == 0.f ? -INFINITY : (ts.m + log2f(ts.s)) / LOG2E;
  const bool ts_only = lse_ts > text_max;   // timestamps carry more mass than any text token: text is masked
  float lse = lse_ts;
  if (!ts_only) {
    MaxSum all = text;
    ms_merge(all, ts.m, ts.s);
    lse = all.s == 0.f ? -INFINITY : (all.m + log2f(all.s)) / LOG2E;
  }
  // ---- merge the 16 warp lists of the chosen kind: warp 0 alone, every lane owns up to MERGE_PER_LANE entries,
  // kc rounds of a warp arg-max (the other 15 warps are done)
  if (wid != 0) return;
  constexpr int MERGE_PER_LANE = (BC_WARPS * BC_KC + 31) / 32;
  float mv[MERGE_PER_LANE];
  int mi[MERGE_PER_LANE];
#pragma unroll
  for (int k = 0; k < MERGE_PER_LANE; ++k) {
    const int ent = lane + 32 * k;
    mv[k] = -INFINITY;
    mi[k] = 0x7fffffff;
    if (ent < BC_WARPS * kc) {
      const WarpList& src = lists[ts_only ? 1 : 0][ent / kc];
      mv[k] = src.v[ent % kc];
      mi[k] = src.i[ent % kc];
    }
  }
  const float cum = p.bm.cum[row];
  for (int r = 0; r < kc; ++r) {
    float bv = mv[0];
    int bi = mi[0];
#pragma unroll
    for (int k = 1; k < MERGE_PER_LANE; ++k)
      if (cand_before(mv[k], mi[k], bv, bi)) { bv = mv[k]; bi = mi[k]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(FULL_MASK, bv, o);
      const int oi = __shfl_xor_sync(FULL_MASK, bi, o);
      if (cand_before(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
#pragma unroll
    for (int k = 0; k < MERGE_PER_LANE; ++k)
      if (bi != 0x7fffffff && mi[k] == bi) { mv[k] = -INFINITY; mi[k] = 0x7fffffff; }   // an index sits in one list only
    if (lane == 0) {
      const bool none = bi == 0x7fffffff;
      p.bm.cand_score[(long long)row * kc + r] = none ? -INFINITY : cum + (bv - lse);
      p.bm.cand_tok[(long long)row * kc + r] = none ? p.eot : bi;
    }
  }
}

// ------------------------------------------------------------------------------------------
// update: one CTA per stream. Thread 0 walks the beam * (beam+1) candidates; all threads then rewrite the
// stream's token histories and indirection rows in place (staged through shared memory).
// ------------------------------------------------------------------------------------------
constexpr int BU_THREADS = 128;
constexpr int BU_NC = JANUS_MAX_BEAM * BC_KC;

__global__ void __launch_bounds__(BU_THREADS) dec_beam_update_kernel(const BeamStepParams p) {
  extern __shared__ __align__(16) uint8_t bu_smem[];
  __shared__ int parent[JANUS_MAX_BEAM], newtok[JANUS_MAX_BEAM];
  __shared__ float newcum[JANUS_MAX_BEAM];
  __shared__ int fin_parent[JANUS_MAX_BEAM];
  __shared__ float fin_sc[JANUS_MAX_BEAM];
  __shared__ int n_newfin, any_move;
  const int b = blockIdx.x, tid = threadIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const int pos = *p.pos, nxt = pos + 1;
  if (nxt >= p.max_len || nxt < p.n_prompt) return;   // prompt feeding: rows already hold the prompt, indir is identity
  const int K = p.bm.beam, kc = K + 1, L = p.max_len, SL = p.self_ld;
  int32_t* toks = p.tokens + (long long)b * K * L;
  if (p.bm.stream_done[b]) {   // finished stream: its rows keep consuming eot
    if (tid < K) toks[(long long)tid * L + nxt] = p.eot;
    return;
  }
  int32_t* hist = reinterpret_cast<int32_t*>(bu_smem);           // [K][nxt]
  uint8_t* ind = bu_smem + (size_t)K * L * sizeof(int32_t);      // [K][SL]
  uint8_t* indir = p.bm.indir + (long long)b * K * SL;
  if (tid == 0) {
    const float* cs = p.bm.cand_score + (long long)b * K * kc;
    const int32_t* ct = p.bm.cand_tok + (long long)b * K * kc;
    float sc[BU_NC];
    int tk[BU_NC];
    const int n = K * kc;
    for (int i = 0; i < n; ++i) { sc[i] = cs[i]; tk[i] = ct[i]; }
    int saved = 0, nf = 0, moved = 0;
    while (saved < K) {
      int best = -1;   // flat order (beam, rank) is (beam, token-sorted-by-value): first strict maximum wins ties by beam;
      for (int i = 0; i < n; ++i) {   // within a beam equal values are already in ascending token order
        if (sc[i] == -INFINITY) continue;
        if (best < 0 || sc[i] > sc[best]) best = i;
      }
      if (best < 0) break;
      const int k = best / kc;
      if (tk[best] == p.eot) {
        if (nf < JANUS_MAX_BEAM) { fin_parent[nf] = k; fin_sc[nf] = sc[best]; ++nf; }
      } else {
        parent[saved] = k; newtok[saved] = tk[best]; newcum[saved] = sc[best];
        moved |= k != saved;
        ++saved;
      }
      sc[best] = -INFINITY;
    }
    for (; saved < K; ++saved) { parent[saved] = saved; newtok[saved] = p.eot; newcum[saved] = -INFINITY; }   // dead beams
    n_newfin = nf;
    any_move = moved;
  }
  __syncthreads();
  const int nf = n_newfin;
  const bool stage = any_move || nf > 0;
  if (stage) {
    for (int i = tid; i < K * nxt; i += BU_THREADS) { const int k = i / nxt, j = i - k * nxt; hist[k * L + j] = toks[(long long)k * L + j]; }
    if (any_move)
      for (int i = tid; i < K * pos; i += BU_THREADS) { const int k = i / pos, j = i - k * pos; ind[k * SL + j] = indir[(long long)k * SL + j]; }
    __syncthreads();
  }
  // newly finished hypotheses, in descending score, while the list has room
  int nfin = p.bm.n_fin[b];
  for (int f = 0; f < nf && nfin < p.bm.max_candidates; ++f, ++nfin) {
    int32_t* dst = p.bm.fin_tokens + ((long long)b * p.bm.max_candidates + nfin) * L;
    const int32_t* src = hist + fin_parent[f] * L;
    for (int j = tid; j < L; j += BU_THREADS) dst[j] = j < nxt ? src[j] : p.eot;
    if (tid == 0) {
      p.bm.fin_score[(long long)b * p.bm.max_candidates + nfin] = fin_sc[f];
      p.bm.fin_len[(long long)b * p.bm.max_candidates + nfin] = nxt - p.n_prompt;
    }
  }
  // next beams
  for (int k = 0; k < K; ++k) {
    const int pk = parent[k];
    if (pk != k) {
      for (int j = tid; j < nxt; j += BU_THREADS) toks[(long long)k * L + j] = hist[pk * L + j];
      for (int j = tid; j < pos; j += BU_THREADS) indir[(long long)k * SL + j] = ind[pk * SL + j];
    }
  }
  if (tid < K) {
    toks[(long long)tid * L + nxt] = newtok[tid];
    indir[(long long)tid * SL + pos] = (uint8_t)parent[tid];   // position `pos` was appended to the parent's slot this step
    p.bm.cum[(long long)b * K + tid] = newcum[tid];
  }
  if (tid == 0) {
    p.bm.n_fin[b] = nfin;
    if (nfin >= p.bm.max_candidates) { p.bm.stream_done[b] = 1; atomicAdd(p.n_done, 1); }
  }
}

// ------------------------------------------------------------------------------------------
// finalize: one CTA per stream
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BU_THREADS)
dec_beam_finalize_kernel(const int32_t* tokens, const int32_t* pos_p, int max_len, int n_prompt, int eot,
                         const janus_beam_state_t bm, int32_t* best_tokens, float* best_score, int32_t* best_len) {
  __shared__ int sel_kind, sel_idx, sel_len;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int K = bm.beam, L = max_len;
  const int n_tok = min(*pos_p + 1, L);        // tokens held by a live row (prompt included)
  if (tid == 0) {
    const int live_len = max(n_tok - n_prompt, 0);
    const int nfin = bm.n_fin[b];
    float best = -INFINITY;
    int kind = -1, idx = 0, len = 0, pool = 0;
    for (int f = 0; f < nfin; ++f, ++pool) {
      const int l = bm.fin_len[(long long)b * bm.max_candidates + f];
      const float v = bm.fin_score[(long long)b * bm.max_candidates + f] / (float)max(l, 1);
      if (kind < 0 || v > best) { best = v; kind = 0; idx = f; len = l; }
    }
    bool used[JANUS_MAX_BEAM];
    for (int k = 0; k < K; ++k) used[k] = false;
    for (; pool < K; ++pool) {   // live beams by descending cumulative score (ties: beam index)
      int bk = -1;
      for (int k = 0; k < K; ++k)
        if (!used[k] && (bk < 0 || bm.cum[(long long)b * K + k] > bm.cum[(long long)b * K + bk])) bk = k;
      if (bk < 0) break;
      used[bk] = true;
      const float v = bm.cum[(long long)b * K + bk] / (float)max(live_len, 1);
      if (kind < 0 || v > best) { best = v; kind = 1; idx = bk; len = live_len; }
    }
    sel_kind = kind; sel_idx = idx; sel_len = len;
    best_score[b] = kind == 0 ? bm.fin_score[(long long)b * bm.max_candidates + idx] : bm.cum[(long long)b * K + idx];
    best_len[b] = len;
  }
  __syncthreads();
  const int32_t* src = sel_kind == 0 ? bm.fin_tokens + ((long long)b * bm.max_candidates + sel_idx) * L
                                     : tokens + ((long long)b * K + sel_idx) * L;
  const int valid = n_prompt + sel_len;
  for (int j = tid; j < L; j += BU_THREADS) best_tokens[(long long)b * L + j] = j < valid ? src[j] : eot;
}

int dec_beam_step(const BeamStepParams& p, cudaStream_t st) {
  const int rows = p.bm.n_streams * p.bm.beam;
  JANUS_CHECK_ARG(p.n_vocab_pad % 4 == 0, "beam: n_vocab_pad must be a multiple of 4");
  const size_t smem = (size_t)p.bm.beam * p.max_len * sizeof(int32_t) + (size_t)p.bm.beam * p.self_ld;
  JANUS_CHECK_ARG(smem <= 40 * 1024, "beam: beam=%d x max_len=%d histories do not fit shared memory", p.bm.beam, p.max_len);
  JANUS_CUDA(launch_pdl(dec_beam_cand_kernel, dim3(rows), dim3(BC_THREADS), 0, st, p));
  JANUS_CUDA(launch_pdl(dec_beam_update_kernel, dim3(p.bm.n_streams), dim3(BU_THREADS), smem, st, p));
  return JANUS_OK;
}

int dec_beam_finalize(const int32_t* tokens, const int32_t* pos, int max_len, int n_prompt, int eot,
                      const janus_beam_state_t& bm, int32_t* best_tokens, float* best_score, int32_t* best_len,
                      cudaStream_t st) {
  dec_beam_finalize_kernel<<<bm.n_streams, BU_THREADS, 0, st>>>(tokens, pos, max_len, n_prompt, eot, bm, best_tokens,
                                                                 best_score, best_len);
  JANUS_LAUNCH_CHECK("dec_beam_finalize_kernel");
  return JANUS_OK;
}

}  // namespace janus
