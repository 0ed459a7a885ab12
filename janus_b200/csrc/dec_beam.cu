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

struct RuleState {   // ApplyTimestampRules / SuppressBlank state of one row (same rules as dec_pick_kernel)
  int n_gen, tb, ts_floor, first_hi, eot, blank;
  bool last_ts, penult_ts;
};

__device__ __forceinline__ RuleState rule_state(const int32_t* toks, int nxt, int n_prompt, int tb, int max_initial_ts,
                                                int eot, int blank) {
  RuleState r;
  r.n_gen = nxt - n_prompt;
  r.tb = tb;
  r.eot = eot;
  r.blank = blank;
  const int last = r.n_gen >= 1 ? toks[nxt - 1] : -1;
  const int penult = r.n_gen >= 2 ? toks[nxt - 2] : -1;
  r.last_ts = r.n_gen >= 1 && last >= tb;
  r.penult_ts = r.n_gen < 2 || penult >= tb;
  r.ts_floor = tb;
  int last_stamp = -1;
  for (int i = n_prompt; i < nxt; ++i) {
    const int t = toks[i];
    if (t >= tb) last_stamp = t;
  }
  if (last_stamp >= 0) r.ts_floor = (r.last_ts && !r.penult_ts) ? last_stamp : last_stamp + 1;
  r.first_hi = tb + max_initial_ts;
  return r;
}

__device__ __forceinline__ bool rule_allows(const RuleState& r, int i, bool suppressed) {
  bool ok = !suppressed;
  if (r.n_gen == 0) ok = ok && i != r.blank && i != r.eot && i >= r.tb && i <= r.first_hi;
  if (r.last_ts) ok = ok && (r.penult_ts ? i < r.tb : i >= r.eot);
  if (i >= r.tb) ok = ok && i >= r.ts_floor;
  return ok;
}

struct MaxSum {   // running max and sum exp(x - max)
  float m, s;
};
__device__ __forceinline__ void ms_add(MaxSum& a, float x) {
  if (x > a.m) {
    a.s = a.s * __expf(a.m - x) + 1.f;
    a.m = x;
  } else {
    a.s += __expf(x - a.m);
  }
}
__device__ __forceinline__ void ms_merge(MaxSum& a, float om, float os) {
  if (om == -INFINITY) return;
  if (a.m == -INFINITY) { a.m = om; a.s = os; return; }
  if (om > a.m) {
    a.s = a.s * __expf(a.m - om) + os;
    a.m = om;
  } else {
    a.s += os * __expf(om - a.m);
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
// candidates: one CTA per row. Pass 1: max / sum-exp of the allowed text and timestamp tokens (decides the
// "timestamps outweigh every text token" rule and the normaliser). Pass 2 (the row is L2 / L1 resident): every
// thread keeps the top kc of its tokens in registers; kc rounds of a block arg-max merge them.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BC_THREADS) dec_beam_cand_kernel(const BeamStepParams p) {
  __shared__ float sh_m[2][BC_THREADS / 32], sh_s[2][BC_THREADS / 32];
  __shared__ float win_v[2][BC_THREADS / 32];
  __shared__ int win_i[2][BC_THREADS / 32];
  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  pdl_launch_dependents();
  pdl_wait();
  const int nxt = *p.pos + 1;
  if (nxt >= p.max_len || nxt < p.n_prompt) return;
  const int K = p.bm.beam, kc = K + 1;
  if (p.bm.stream_done[row / K]) return;
  const int32_t* toks = p.tokens + (long long)row * p.max_len;
  const RuleState rs = rule_state(toks, nxt, p.n_prompt, p.ts_begin, p.max_initial_ts, p.eot, p.blank);
  const float* lg = p.logits + (long long)row * p.n_vocab_pad;
  const int tb = p.ts_begin;
  // ---- pass 1
  MaxSum text = {-INFINITY, 0.f}, ts = {-INFINITY, 0.f};
  for (int i0 = tid * 4; i0 < p.n_vocab; i0 += BC_THREADS * 4) {
    const float4 x4 = *reinterpret_cast<const float4*>(lg + i0);
    const uint32_t sup4 = load_suppress4(p.suppress, i0, p.n_vocab);
    const float xs[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int i = i0 + e;
      if (!rule_allows(rs, i, ((sup4 >> (8 * e)) & 0xffu) != 0)) continue;
      if (i < tb) ms_add(text, xs[e]); else ms_add(ts, xs[e]);
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
  for (int w = 1; w < BC_THREADS / 32; ++w) { ms_merge(text, sh_m[0][w], sh_s[0][w]); ms_merge(ts, sh_m[1][w], sh_s[1][w]); }
  const float lse_ts = ts.m == -INFINITY ? -INFINITY : ts.m + logf(ts.s);
  const bool ts_only = lse_ts > text.m;   // timestamps carry more mass than any text token: text is masked
  float lse = lse_ts;
  if (!ts_only) {
    MaxSum all = text;
    ms_merge(all, ts.m, ts.s);
    lse = all.m == -INFINITY ? -INFINITY : all.m + logf(all.s);
  }
  // ---- pass 2: per-thread top kc (sorted, registers)
  float lv[BC_KC];
  int li[BC_KC];
#pragma unroll
  for (int j = 0; j < BC_KC; ++j) { lv[j] = -INFINITY; li[j] = 0x7fffffff; }
  for (int i0 = tid * 4; i0 < p.n_vocab; i0 += BC_THREADS * 4) {
    const float4 x4 = *reinterpret_cast<const float4*>(lg + i0);
    const uint32_t sup4 = load_suppress4(p.suppress, i0, p.n_vocab);
    const float xs[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int i = i0 + e;
      const float x = xs[e];
      if (!rule_allows(rs, i, ((sup4 >> (8 * e)) & 0xffu) != 0) || (ts_only && i < tb)) continue;
      bool gt_last = false;   // x beats the thread's current kc-th value (indices ascend, so ties stay behind)
#pragma unroll
      for (int j = 0; j < BC_KC; ++j)
        if (j == kc - 1) gt_last = x > lv[j];
      if (!gt_last) continue;
#pragma unroll
      for (int j = BC_KC - 1; j >= 0; --j) {
        if (j < kc) {
          if (j > 0 && x > lv[j - 1]) { lv[j] = lv[j - 1]; li[j] = li[j - 1]; }
          else if (x > lv[j]) { lv[j] = x; li[j] = i; }
        }
      }
    }
  }
  // ---- kc rounds of block arg-max over the threads' list heads
  const float cum = p.bm.cum[row];
  int head = 0;
  for (int r = 0; r < kc; ++r) {
    float v = -INFINITY;
    int idx = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < BC_KC; ++j)
      if (j == head) { v = lv[j]; idx = li[j]; }
    float bv = v;
    int bi = idx;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(FULL_MASK, bv, o);
      const int oi = __shfl_xor_sync(FULL_MASK, bi, o);
      if (cand_before(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) { win_v[r & 1][wid] = bv; win_i[r & 1][wid] = bi; }
    __syncthreads();
    bv = win_v[r & 1][0]; bi = win_i[r & 1][0];
    for (int w = 1; w < BC_THREADS / 32; ++w)
      if (cand_before(win_v[r & 1][w], win_i[r & 1][w], bv, bi)) { bv = win_v[r & 1][w]; bi = win_i[r & 1][w]; }
    if (bi != 0x7fffffff && bi == idx) ++head;   // every token index belongs to exactly one thread
    if (tid == 0) {
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
