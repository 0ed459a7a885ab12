/*
 * oracle/yin_aubio.c -- TEST INFRASTRUCTURE ONLY (CPU oracle; never on the product path).
 *
 * CPU restatement, in strict sequential float32, of the arithmetic that the
 * reference's ProsodyExtractor delegates to the third-party package `aubio`
 * (pinned only as `aubio>=0.4.9`, /root/reference/backend/requirements.txt:4;
 * call sites /root/reference/backend/services/prosody.py:32-34 (constructor:
 * method 'yin', buf 4096, hop 512, unit 'Hz', tolerance 0.8) and :84 (one call
 * per 512-sample hop)).
 *
 * aubio is NOT vendored in the reference and is not installable in this
 * environment, so this file restates aubio 0.4.9's published algorithm:
 *   src/pitch/pitch.c      aubio_pitch_do, aubio_pitch_slideblock, aubio_pitch_do_yin
 *   src/pitch/pitchyin.c   aubio_pitchyin_do (difference fn + cumulative mean
 *                          normalisation + early exit at period = tau-3)
 *   src/mathutils.c        fvec_quadratic_peak_pos, fvec_min_elem,
 *                          aubio_level_lin, aubio_db_spl, aubio_silence_detection
 * smpl_t is float (aubio's default single-precision build, as shipped in the
 * PyPI wheels). Build with -ffp-contract=off so that no FMA is formed, which is
 * what an x86-64 baseline wheel executes.
 *
 * PARITY UNPINNED: the reference's own tests only pin tag-level facts for this
 * path (test_input_processing.py:447-505); per-frame f0 has no golden vectors.
 *
 * The oracle also tallies the algorithmic work actually performed (number of
 * lags evaluated before the early exit) so that bench.py can credit early-exit
 * work exactly.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef float smpl_t;

typedef struct {
  uint32_t buf_size;    /* aubio bufsize (4096 in the reference)            */
  uint32_t hop_size;    /* 512                                              */
  uint32_t samplerate;  /* 48000                                            */
  smpl_t tol;           /* 0.8 (reference) ; aubio default 0.15             */
  smpl_t silence_db;    /* aubio DEFAULT_PITCH_SILENCE = -50                */
  smpl_t *buf;          /* sliding buffer, persists across calls (stateful) */
  smpl_t *yin;          /* buf_size/2 lags                                  */
  uint64_t lags_evaluated; /* work tally: sum over hops of (tau_exit + 1)   */
} yin_oracle_t;

yin_oracle_t *yin_oracle_new(uint32_t buf_size, uint32_t hop_size, uint32_t samplerate,
                             float tol, float silence_db) {
  yin_oracle_t *o = (yin_oracle_t *)calloc(1, sizeof(*o));
  if (!o) return NULL;
  o->buf_size = buf_size;
  o->hop_size = hop_size;
  o->samplerate = samplerate;
  o->tol = tol;
  o->silence_db = silence_db;
  o->buf = (smpl_t *)calloc(buf_size, sizeof(smpl_t));
  o->yin = (smpl_t *)calloc(buf_size / 2, sizeof(smpl_t));
  return o;
}

void yin_oracle_del(yin_oracle_t *o) {
  if (!o) return;
  free(o->buf);
  free(o->yin);
  free(o);
}

void yin_oracle_reset(yin_oracle_t *o) {
  memset(o->buf, 0, o->buf_size * sizeof(smpl_t));
  memset(o->yin, 0, (o->buf_size / 2) * sizeof(smpl_t));
  o->lags_evaluated = 0;
}

void yin_oracle_get_state(const yin_oracle_t *o, float *out) {
  memcpy(out, o->buf, o->buf_size * sizeof(smpl_t));
}

void yin_oracle_set_state(yin_oracle_t *o, const float *in) {
  memcpy(o->buf, in, o->buf_size * sizeof(smpl_t));
}

uint64_t yin_oracle_lags(const yin_oracle_t *o) { return o->lags_evaluated; }

/* mathutils.c: fvec_quadratic_peak_pos */
static smpl_t quadratic_peak_pos(const smpl_t *x, uint32_t length, uint32_t pos) {
  smpl_t s0, s1, s2;
  uint32_t x0, x2;
  smpl_t half = .5f, two = 2.f;
  if (pos == 0 || pos == length - 1) return (smpl_t)pos;
  x0 = (pos < 1) ? pos : pos - 1;
  x2 = (pos + 1 < length) ? pos + 1 : pos;
  if (x0 == pos) return (x[pos] <= x[x2]) ? (smpl_t)pos : (smpl_t)x2;
  if (x2 == pos) return (x[pos] <= x[x0]) ? (smpl_t)pos : (smpl_t)x0;
  s0 = x[x0];
  s1 = x[pos];
  s2 = x[x2];
  return pos + half * (s0 - s2) / (s0 - two * s1 + s2);
}

/* mathutils.c: fvec_min_elem -- note: on ties the LAST index wins */
static uint32_t min_elem(const smpl_t *s, uint32_t length) {
  uint32_t j, pos = 0;
  smpl_t tmp = s[0];
  for (j = 0; j < length; j++) {
    pos = (tmp < s[j]) ? pos : j;
    tmp = (tmp < s[j]) ? tmp : s[j];
  }
  return pos;
}

/* pitchyin.c: aubio_pitchyin_do. Returns the (fractional) period in samples. */
static smpl_t pitchyin_do(yin_oracle_t *o, const smpl_t *input) {
  const smpl_t tol = o->tol;
  smpl_t *yin = o->yin;
  const uint32_t length = o->buf_size / 2;
  uint32_t j, tau;
  int32_t period;
  smpl_t tmp, tmp2 = 0.f;

  yin[0] = 1.f;
  for (tau = 1; tau < length; tau++) {
    yin[tau] = 0.f;
    for (j = 0; j < length; j++) {
      tmp = input[j] - input[j + tau];
      yin[tau] += tmp * tmp;
    }
    tmp2 += yin[tau];
    if (tmp2 != 0) {
      yin[tau] *= tau / tmp2;
    } else {
      yin[tau] = 1.f;
    }
    period = (int32_t)tau - 3;
    if (tau > 4 && (yin[period] < tol) && (yin[period] < yin[period + 1])) {
      o->lags_evaluated += tau + 1;
      return quadratic_peak_pos(yin, length, (uint32_t)period);
    }
  }
  o->lags_evaluated += length;
  return quadratic_peak_pos(yin, length, min_elem(yin, length));
}

/* mathutils.c: aubio_level_lin / aubio_db_spl on one hop */
static smpl_t level_lin(const smpl_t *f, uint32_t length) {
  smpl_t energy = 0.f;
  uint32_t j;
  for (j = 0; j < length; j++) energy += f[j] * f[j];
  return energy / length;
}

/*
 * One aubio.pitch('yin', ...) __call__ on one hop (pitch.c: aubio_pitch_do ->
 * aubio_pitch_do_yin -> slideblock + pitchyin_do; silence gate on the hop;
 * unit 'Hz' conversion is the identity).
 * Also returns the hop's level in dB (10*log10(mean(x^2))) through *db_out.
 */
float yin_oracle_do(yin_oracle_t *o, const float *hop, float *db_out) {
  const uint32_t overlap = o->buf_size - o->hop_size;
  uint32_t j;
  smpl_t pitch, db;
  /* aubio_pitch_slideblock */
  for (j = 0; j < overlap; j++) o->buf[j] = o->buf[j + o->hop_size];
  for (j = 0; j < o->hop_size; j++) o->buf[j + overlap] = hop[j];
  /* aubio_pitch_do_yin */
  pitch = pitchyin_do(o, o->buf);
  if (pitch > 0) {
    pitch = o->samplerate / (pitch + 0.); /* double division, as in aubio */
  } else {
    pitch = 0.f;
  }
  /* aubio_pitch_do: silence gate on the incoming hop */
  db = 10.f * log10f(level_lin(hop, o->hop_size));
  if (db_out) *db_out = db;
  if (db < o->silence_db) pitch = 0.f;
  return pitch;
}

/*
 * Convenience driver: runs the reference's hop loop
 * (/root/reference/backend/services/prosody.py:76-87) over a whole buffer:
 * ceil(n/hop) hops, last hop zero-padded. f0_out/db_out have ceil(n/hop) entries.
 */
void yin_oracle_run(yin_oracle_t *o, const float *pcm, uint64_t n, float *f0_out, float *db_out) {
  const uint32_t hop = o->hop_size;
  smpl_t *chunk = (smpl_t *)malloc(hop * sizeof(smpl_t));
  uint64_t i, h = 0;
  for (i = 0; i < n; i += hop, h++) {
    uint64_t m = (n - i < hop) ? (n - i) : hop;
    const float *src = pcm + i;
    if (m < hop) {
      memcpy(chunk, src, m * sizeof(smpl_t));
      memset(chunk + m, 0, (hop - m) * sizeof(smpl_t));
      src = chunk;
    }
    f0_out[h] = yin_oracle_do(o, src, db_out ? db_out + h : NULL);
  }
  free(chunk);
}

/*
 * numpy's float32 pairwise summation (numpy/_core/src/umath/loops_utils.h.src,
 * @TYPE@_pairwise_sum), which is what np.mean(list_of_float32) executes at
 * /root/reference/backend/services/prosody.py:90. Restated here so the GPU's
 * per-stream mean-f0 reduction can be checked bit-for-bit.
 */
static float np_pairwise_sum_f32(const float *a, int64_t n) {
  if (n < 8) {
    int64_t i;
    float res = 0.f;
    for (i = 0; i < n; i++) res += a[i];
    return res;
  } else if (n <= 128) {
    int64_t i;
    float r[8], res;
    for (i = 0; i < 8; i++) r[i] = a[i];
    for (i = 8; i < n - (n % 8); i += 8) {
      r[0] += a[i + 0]; r[1] += a[i + 1]; r[2] += a[i + 2]; r[3] += a[i + 3];
      r[4] += a[i + 4]; r[5] += a[i + 5]; r[6] += a[i + 6]; r[7] += a[i + 7];
    }
    res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < n; i++) res += a[i];
    return res;
  } else {
    int64_t n2 = n / 2;
    n2 -= n2 % 8;
    return np_pairwise_sum_f32(a, n2) + np_pairwise_sum_f32(a + n2, n - n2);
  }
}

float np_mean_f32(const float *a, int64_t n) { return np_pairwise_sum_f32(a, n) / (float)n; }
