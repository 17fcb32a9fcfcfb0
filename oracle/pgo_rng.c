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
double pgo_pcg64_double(pgo_pcg64* g) { return (double)(pgo_pcg64_next(g) >> 11) * (1.0 / 9007199254740992.0); }
double pgo_pcg64_uniform(pgo_pcg64* g, double lo, double hi) { return lo + (hi - lo) * pgo_pcg64_double(g); }
