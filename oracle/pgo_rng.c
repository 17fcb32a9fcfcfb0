/* oracle/pgo_rng.c -- numpy.random.Generator(PCG64(SeedSequence(seed))) restated, as used by
 * gymnasium.utils.seeding.np_random(seed) (reference core.py:302: task.np_random is re-created per reset).
 * TEST INFRASTRUCTURE ONLY.  Checked bit-for-bit against numpy in tests/test_oracle_rng.py.
 */
#include "pgo.h"

typedef unsigned __int128 u128;

static uint32_t hashmix(uint32_t value, uint32_t* hc) {
  value ^= *hc; *hc *= 0x931e8875u; value *= *hc; value ^= value >> 16; return value;
}
static uint32_t mix(uint32_t x, uint32_t y) {
  uint32_t r = 0xca01f9ddu * x - 0x4973f715u * y; r ^= r >> 16; return r;
}

void pgo_pcg64_seed(pgo_pcg64* g, uint64_t seed) {
  /* SeedSequence(entropy=seed, pool_size=4) */
  uint32_t ent[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  int nent = ent[1] ? 2 : 1;
  uint32_t pool[4], hc = 0x43b0d7e5u;
  for (int i = 0; i < 4; i++) pool[i] = hashmix(i < nent ? ent[i] : 0u, &hc);
  for (int s = 0; s < 4; s++)
    for (int d = 0; d < 4; d++)
      if (s != d) pool[d] = mix(pool[d], hashmix(pool[s], &hc));
  /* generate_state(4, uint64) == 8 uint32 words, low word first */
  uint32_t w[8], hb = 0x8b51f9ddu;
  for (int i = 0; i < 8; i++) {
    uint32_t v = pool[i & 3];
    v ^= hb; hb *= 0x58f38dedu; v *= hb; v ^= v >> 16; w[i] = v;
  }
  uint64_t st[4];
  for (int i = 0; i < 4; i++) st[i] = (uint64_t)w[2 * i] | ((uint64_t)w[2 * i + 1] << 32);
  /* pcg64_set_seed(seed = (st0 hi, st1 lo), inc = (st2 hi, st3 lo)); pcg_setseq_128_srandom_r */
  const u128 MULT = ((u128)0x2360ed051fc65da4ULL << 64) | 0x4385df649fccf645ULL;
  u128 initstate = ((u128)st[0] << 64) | st[1], initseq = ((u128)st[2] << 64) | st[3];
  u128 inc = (initseq << 1) | 1, state = 0;
  state = state * MULT + inc;
  state += initstate;
  state = state * MULT + inc;
  g->s_hi = (uint64_t)(state >> 64); g->s_lo = (uint64_t)state;
  g->inc_hi = (uint64_t)(inc >> 64); g->inc_lo = (uint64_t)inc;
}

uint64_t pgo_pcg64_next(pgo_pcg64* g) {
  const u128 MULT = ((u128)0x2360ed051fc65da4ULL << 64) | 0x4385df649fccf645ULL;
  u128 state = ((u128)g->s_hi << 64) | g->s_lo, inc = ((u128)g->inc_hi << 64) | g->inc_lo;
  state = state * MULT + inc;
  g->s_hi = (uint64_t)(state >> 64); g->s_lo = (uint64_t)state;
  uint64_t x = g->s_hi ^ g->s_lo;
  unsigned rot = (unsigned)(g->s_hi >> 58);
  return (x >> rot) | (x << ((-rot) & 63));
}
/* BitGenerator.advance(delta): the state after delta draws, by LCG jump-ahead (Brown, "Random number generation with arbitrary
 * strides") in O(log delta) 128-bit multiplications */
void pgo_pcg64_advance(pgo_pcg64* g, uint64_t delta) {
  const u128 MULT = ((u128)0x2360ed051fc65da4ULL << 64) | 0x4385df649fccf645ULL;
  u128 state = ((u128)g->s_hi << 64) | g->s_lo, inc = ((u128)g->inc_hi << 64) | g->inc_lo;
  u128 acc_mult = 1, acc_plus = 0, cur_mult = MULT, cur_plus = inc;
  while (delta > 0) {
    if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
    cur_plus = (cur_mult + 1) * cur_plus;
    cur_mult *= cur_mult;
    delta >>= 1;
  }
  state = acc_mult * state + acc_plus;
  g->s_hi = (uint64_t)(state >> 64); g->s_lo = (uint64_t)state;
}
double pgo_pcg64_double(pgo_pcg64* g) { return (double)(pgo_pcg64_next(g) >> 11) * (1.0 / 9007199254740992.0); }
double pgo_pcg64_uniform(pgo_pcg64* g, double lo, double hi) { return lo + (hi - lo) * pgo_pcg64_double(g); }

/* Generator.standard_exponential(): numpy's 256-layer ziggurat (random_standard_exponential); tables recovered from and
 * verified against numpy itself by tools/gen_ziggurat.py */
#include "pgo_ziggurat.h"
#include <math.h>
double pgo_pcg64_standard_exponential(pgo_pcg64* g) {
  for (;;) {
    uint64_t ri = pgo_pcg64_next(g) >> 3;
    const int idx = (int)(ri & 0xff);
    ri >>= 8;
    const double x = (double)ri * ZIG_we[idx];
    if (ri < ZIG_ke[idx]) return x;
    const double u = pgo_pcg64_double(g);
    if (idx == 0) return ZIG_R - log1p(-u);
    if ((ZIG_fe[idx - 1] - ZIG_fe[idx]) * u + ZIG_fe[idx] < exp(-x)) return x;
  }
}
/* Generator.dirichlet(alpha=[1, 1, 1]): alpha.max() >= 0.1, so three standard_gamma(1.0) = standard_exponential draws,
 * normalised by the reciprocal of their running sum */
void pgo_pcg64_dirichlet111(pgo_pcg64* g, double out[3]) {
  double acc = 0.0;
  for (int j = 0; j < 3; j++) { out[j] = pgo_pcg64_standard_exponential(g); acc = acc + out[j]; }
  const double inv = 1 / acc;
  for (int j = 0; j < 3; j++) out[j] = out[j] * inv;
}
