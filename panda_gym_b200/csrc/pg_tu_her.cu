// pg_tu_her.cu -- kernels of the on-device HER replay buffer (contract: pg_her.cuh)
#include <cstdio>
#include "pg_her.cuh"
#include "pg_launch.h"

namespace pg {

// ------------------------------------------------------------------------------------------ add
// DictReplayBuffer.add + HerReplayBuffer episode bookkeeping; one thread per environment
__global__ void k_her_add_book(HerParams H, const uint8_t* __restrict__ done) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= H.n_envs) return;
  const int pos = H.cursor[0], rows = H.rows, ne = H.n_envs;
  // overwriting the first transition of an old episode invalidates the rest of that episode
  int st = H.ep_start[(size_t)pos * ne + e], len = H.ep_length[(size_t)pos * ne + e];
  if (len > 0) for (int i = pos; i < st + len; i++) H.ep_length[(size_t)(i % rows) * ne + e] = 0;
  H.ep_start[(size_t)pos * ne + e] = H.cur_ep_start[e];
  H.rec[((size_t)pos * ne + e) * H.rec_w + her_off_eps(H)] = __int_as_float(H.cur_ep_start[e]);
  if (done[e]) {  // _compute_episode_length, evaluated with the cursor already advanced
    int start = H.cur_ep_start[e], end = (pos + 1) % rows;
    if (end < start) end += rows;
    for (int i = start; i < end; i++) H.ep_length[(size_t)(i % rows) * ne + e] = end - start;
    H.cur_ep_start[e] = (pos + 1) % rows;
  }
}
__global__ void k_her_add_data(HerParams H, const float* __restrict__ obs, const float* __restrict__ ag, const float* __restrict__ dg,
                               const float* __restrict__ nobs, const float* __restrict__ nag, const float* __restrict__ ndg,
                               const float* __restrict__ act, const float* __restrict__ rew, const uint8_t* __restrict__ done,
                               const uint8_t* __restrict__ timeout) {
  const int pos = H.cursor[0], ne = H.n_envs, od = H.obs_dim, ad = H.act_dim;
  const int per = 2 * od + 12 + ad + 3;  // everything but ep_start (k_her_add_book writes that slot)
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)ne * per) return;
  int e = (int)(i / per), k = (int)(i % per);
  float* dst = H.rec + ((size_t)pos * ne + e) * H.rec_w + k;  // consecutive threads -> consecutive floats of the record
  float v;
  if (k < od) v = obs[(size_t)e * od + k];
  else if ((k -= od) < od) v = nobs[(size_t)e * od + k];
  else if ((k -= od) < 3) v = ag[(size_t)e * 3 + k];
  else if ((k -= 3) < 3) v = dg[(size_t)e * 3 + k];
  else if ((k -= 3) < 3) v = nag[(size_t)e * 3 + k];
  else if ((k -= 3) < 3) v = ndg[(size_t)e * 3 + k];
  else if ((k -= 3) < ad) v = act[(size_t)e * ad + k];
  else if ((k -= ad) == 0) v = rew[e];
  else if (k == 1) v = done[e] ? 1.f : 0.f;
  else v = timeout ? (timeout[e] ? 1.f : 0.f) : 0.f;
  *dst = v;
}
__global__ void k_her_advance(HerParams H) {
  int pos = H.cursor[0] + 1;
  if (pos == H.rows) { H.cursor[1] = 1; pos = 0; }
  H.cursor[0] = pos;
}

// ------------------------------------------------------------------------------------------ valid-index compaction
__global__ void k_her_count(HerParams H, int total) {
  __shared__ int s[kHerBlock];
  int i = blockIdx.x * kHerBlock + threadIdx.x;
  s[threadIdx.x] = (i < total && H.ep_length[i] > 0) ? 1 : 0;
  __syncthreads();
  for (int o = kHerBlock / 2; o > 0; o >>= 1) { if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) H.block_counts[blockIdx.x] = s[0];
}
__global__ void k_her_scan_blocks(HerParams H, int nblocks) {  // single block: exclusive scan of block_counts
  __shared__ int carry;
  __shared__ int s[1024];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nblocks; base += 1024) {
    int i = base + threadIdx.x;
    int v = i < nblocks ? H.block_counts[i] : 0;
    s[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) { int t = threadIdx.x >= o ? s[threadIdx.x - o] : 0; __syncthreads(); s[threadIdx.x] += t; __syncthreads(); }
    if (i < nblocks) H.block_counts[i] = carry + s[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry += s[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) H.n_valid[0] = carry;
}
__global__ void k_her_scatter(HerParams H, int total) {
  __shared__ int s[kHerBlock];
  int i = blockIdx.x * kHerBlock + threadIdx.x;
  int f = (i < total && H.ep_length[i] > 0) ? 1 : 0;
  s[threadIdx.x] = f;
  __syncthreads();
  for (int o = 1; o < kHerBlock; o <<= 1) { int t = threadIdx.x >= o ? s[threadIdx.x - o] : 0; __syncthreads(); s[threadIdx.x] += t; __syncthreads(); }
  if (f) H.valid[H.block_counts[blockIdx.x] + s[threadIdx.x] - 1] = i;  // row-major order == np.flatnonzero
}

// ------------------------------------------------------------------------------------------ MT19937 (numpy legacy RandomState)
__device__ __forceinline__ uint32_t mt_mix(uint32_t a, uint32_t b) {
  uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
  return (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
}
__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
  y ^= (y >> 11); y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= (y >> 18); return y;
}
// one block regenerates the 624-word state in place (three dependency-free phases + the last word)
__device__ void mt_twist(uint32_t* mt) {
  const int t = threadIdx.x, nt = blockDim.x;
  uint32_t v[3]; int cnt;
  cnt = 0; for (int k = t; k < 227; k += nt) v[cnt++] = mt[k + 397] ^ mt_mix(mt[k], mt[k + 1]);
  __syncthreads();
  cnt = 0; for (int k = t; k < 227; k += nt) mt[k] = v[cnt++];
  __syncthreads();
  cnt = 0; for (int k = 227 + t; k < 454; k += nt) v[cnt++] = mt[k - 227] ^ mt_mix(mt[k], mt[k + 1]);
  __syncthreads();
  cnt = 0; for (int k = 227 + t; k < 454; k += nt) mt[k] = v[cnt++];
  __syncthreads();
  cnt = 0; for (int k = 454 + t; k < 623; k += nt) v[cnt++] = mt[k - 227] ^ mt_mix(mt[k], mt[k + 1]);
  __syncthreads();
  cnt = 0; for (int k = 454 + t; k < 623; k += nt) mt[k] = v[cnt++];
  __syncthreads();
  if (t == 0) mt[623] = mt[396] ^ mt_mix(mt[623], mt[0]);
  __syncthreads();
}

// inclusive prefix sum of one int per thread over a 256-thread block: warp shuffles + one shared exchange of the 8 warp totals
// (two barriers; s_w is reusable after the call returns); *total = the block's sum
__device__ __forceinline__ int block_scan256(int v, int* s_w, int* total) {
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
  if (lane == 31) s_w[w] = x;
  __syncthreads();
  int before = 0, all = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) { const int c = s_w[i]; all += c; if (i < w) before += c; }
  __syncthreads();
  *total = all;
  return x + before;
}

constexpr int kHerChunk = 1024;  // phase-B elements staged in shared memory at a time
// Index sampling, numpy-exact.  ONE block of 256 threads (blockDim >= 128 so each twist phase fits 3 values/thread).
//   phase A: np.random.choice(valid, B) == randint(0, n_valid, B): masked rejection on 32-bit draws, all elements
//            share one mask -> accept flags + ordered compaction, chunk by chunk;
//   phase B: np.random.randint(cur[], ep_len[]) for the first nb_virtual samples: per-element mask, elements in
//            order, an element with range 0 consumes no draw -> a sequential automaton over the raw stream.
// Raw words are produced on demand; the MT state left behind equals numpy's after the same two calls.
__global__ void __launch_bounds__(256) k_her_indices_exact(HerParams H, int B, int nb_virtual) {
  __shared__ uint32_t mt[625];
  __shared__ int s_scan[256];
  __shared__ int s_count, s_stop, s_pos, s_tw_ok;
  __shared__ uint32_t s_rng[kHerChunk], s_tw[624];  // per-element range -> accepted draw; tempered words of the state
  __shared__ int s_cur[kHerChunk], s_st[kHerChunk];
  __shared__ uint32_t s_msk[kHerChunk];   // per-element rejection mask
  __shared__ int s_nzc[kHerChunk + 1];    // exclusive count of non-zero ranges before each element of the chunk
  const int t = threadIdx.x;
  for (int k = t; k < 625; k += blockDim.x) mt[k] = H.mt[k];
  __syncthreads();
  const int nv = H.n_valid[0];
  if (nv <= 0) { if (t == 0) H.status[0] = 1; return; }
  if (t == 0) { H.status[0] = 0; s_count = 0; s_pos = (int)mt[624]; }
  __syncthreads();
  const int rows = H.rows, ne = H.n_envs;
#ifdef PG_HER_TIMING
  long long tA0 = clock64(), tStage = 0, tWalk = 0, tTwist = 0, tc, tLoop = 0; int nRounds = 0;
#endif
  // ---- phase A
  const uint32_t rng = (uint32_t)(nv - 1);
  uint32_t mask = rng; mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
  if (rng == 0) {  // a single valid transition: no draws
    for (int k = t; k < B; k += blockDim.x) { int idx = H.valid[0]; H.s_row[k] = idx / ne; H.s_env[k] = idx % ne; }
    __syncthreads();
  } else {
    for (;;) {
      if (s_pos >= 624) { __syncthreads(); mt_twist(mt); if (t == 0) s_pos = 0; __syncthreads(); }
      int pos = s_pos, avail = 624 - pos, chunk = avail < 256 ? avail : 256;
      uint32_t w = 0; int f = 0;
      if (t < chunk) { w = mt_temper(mt[pos + t]) & mask; f = w <= rng; }
      int tot;
      const int incl = block_scan256(f, s_scan, &tot);
      int before = s_count, mine = before + incl - 1;  // index of this accepted draw
      if (t == 0) s_stop = -1;
      __syncthreads();
      PG_CHECK(!(f && mine < B) || (mine >= 0 && w <= rng));
      if (f && mine < B) { H.s_row[mine] = (int)w; if (mine == B - 1) s_stop = t; }  // the raw draw; resolved below, off this loop's critical path
      __syncthreads();
      if (t == 0) {
        if (s_stop >= 0) { s_pos = pos + s_stop + 1; s_count = B; }
        else { s_pos = pos + chunk; s_count = before + tot; }
      }
      __syncthreads();
      if (s_count >= B) break;
    }
    // draw -> transition: valid[] is a random gather over megabytes (a DRAM round trip each); done for all B at once, the
    // loads of a thread independent of each other, instead of one dependent load per 256-word iteration above
    __syncthreads();
    for (int k = t; k < B; k += blockDim.x) { const int idx = H.valid[H.s_row[k]]; H.s_row[k] = idx / ne; H.s_env[k] = idx % ne; }
    __syncthreads();
  }
  // ---- phase B (future strategy): t = randint(cur, ep_len) per virtual sample.  numpy draws element by element with a
  // PER-ELEMENT mask and an element whose range is 0 consumes nothing: a sequential automaton over the raw stream.  Only
  // that automaton is serial: the per-element ranges (dependent global-memory gathers) and the tempering of the 624
  // state words are done by the whole block, chunk by chunk; thread 0 walks shared memory only (a few ns per element).
  // The state is regenerated lazily, when a draw is needed and none is left, exactly like numpy.
  __syncthreads();
#ifdef PG_HER_TIMING
  long long tA1 = clock64();
#endif
  if (t == 0) s_tw_ok = 0;
  for (int c0 = 0; c0 < nb_virtual; c0 += kHerChunk) {
#ifdef PG_HER_TIMING
    tc = clock64();
#endif
    const int cn = nb_virtual - c0 < kHerChunk ? nb_virtual - c0 : kHerChunk;
    __syncthreads();
    // staging, whole block: ranges, masks and the running count of elements that will draw at all.  Thread t owns the 4
    // consecutive elements 4t .. 4t+3 so that the prefix count is one block scan of per-thread totals.
    {
      int nzl = 0;
#pragma unroll
      for (int j = 0; j < kHerChunk / 256; j++) {
        const int k = t * (kHerChunk / 256) + j;
        uint32_t r = 0;
        if (k < cn) {
          const int row = H.s_row[c0 + k], env = H.s_env[c0 + k];
          const int st = H.ep_start[(size_t)row * ne + env], len = H.ep_length[(size_t)row * ne + env];
          const int cur = ((row - st) % rows + rows) % rows;
          r = (uint32_t)(len - 1 - cur);  // high - 1 - low
          s_cur[k] = cur; s_st[k] = st;
        }
        uint32_t m = r; m |= m >> 1; m |= m >> 2; m |= m >> 4; m |= m >> 8; m |= m >> 16;
        s_rng[k] = r; s_msk[k] = m;
        nzl += r != 0u;
      }
      int tot;
      int run = block_scan256(nzl, s_scan, &tot) - nzl;  // exclusive prefix of this thread's first element
#pragma unroll
      for (int j = 0; j < kHerChunk / 256; j++) {
        const int k = t * (kHerChunk / 256) + j;
        s_nzc[k] = run;
        run += s_rng[k] != 0u;
      }
      if (t == 255) s_nzc[kHerChunk] = run;
    }
    if (t == 0) s_count = 0;
    __syncthreads();
#ifdef PG_HER_TIMING
    tStage += clock64() - tc;
#endif
    while (s_count < cn) {
#ifdef PG_HER_TIMING
      tc = clock64();
#endif
      if (!s_tw_ok) {
        __syncthreads();
        for (int k = t; k < 624; k += blockDim.x) s_tw[k] = mt_temper(mt[k]);
        if (t == 0) s_tw_ok = 1;
        __syncthreads();
      }
      if (t < 32) {
        // Warp 0 walks the automaton SPECULATIVELY, 32 elements per round.  Lane i takes element k + i and assumes that no
        // element before it rejects a word: its first word is then pos + (non-zero ranges among k .. k+i-1), a difference
        // of two staged prefix counts.  It tests that word and the two after it (shift 0, 1, 2).  Up to the first lane
        // that rejects at shift 0 the assumption held; from that lane on everybody is one word late, so the round goes on
        // at shift 1 up to the next rejection, then at shift 2.  A round so commits three runs of accepting elements
        // (acceptance probability (r + 1) / (mask + 1) >= 1/2) for roughly the latency of one element of a sequential walk.
        int pos = s_pos, k = s_count, need = 0;
#ifdef PG_HER_TIMING
        long long tw0 = clock64();
#endif
        while (k < cn) {
#ifdef PG_HER_TIMING
          nRounds++;
#endif
          const int idx = k + t;
          const bool valid = idx < cn;
          const int ic = valid ? idx : cn;
          const uint32_t r = valid ? s_rng[idx] : 0u, m = valid ? s_msk[idx] : 0u;
          PG_CHECK(k >= 0 && k < kHerChunk && ic >= 0 && ic <= kHerChunk && pos >= 0 && pos <= 624);
          const int p0 = pos + (s_nzc[ic] - s_nzc[k]);
          PG_CHECK(p0 >= pos);
          const bool nz = r != 0u;
          uint32_t v[3]; unsigned F[3], I[3];
#pragma unroll
          for (int sft = 0; sft < 3; sft++) {
            const bool in = p0 + sft < 624;
            v[sft] = (nz && in) ? (s_tw[p0 + sft] & m) : 0u;
            F[sft] = __ballot_sync(0xffffffffu, valid && nz && !(in && v[sft] <= r));
            I[sft] = __ballot_sync(0xffffffffu, in);
          }
          const unsigned nzm = __ballot_sync(0xffffffffu, nz);
          const int nvalid = __popc(__ballot_sync(0xffffffffu, valid));
          int cur = 0, rej = 0;
#pragma unroll
          for (int sft = 0; sft < 3; sft++) {
            const unsigned Fs = F[sft] & ~((1u << cur) - 1u);  // rejections among the lanes not committed yet (cur < 32 here)
            const int f = Fs ? __ffs(Fs) - 1 : nvalid;
            PG_CHECK(!(t >= cur && t < f) || (idx >= 0 && idx < cn));
            if (t >= cur && t < f) s_rng[idx] = v[sft];          // the slot now holds the accepted draw (0 for a zero range)
            cur = f;
            if (!Fs) break;
            if (!((I[sft] >> f) & 1u)) { need = 1; break; }      // no word left for element k + f: regenerate, retry it
            rej++;                                               // the rejected word is consumed; element k + f tries the next one
          }
          pos += __popc(nzm & ((cur >= 32) ? 0xffffffffu : ((1u << cur) - 1u))) + rej;
          k += cur;
          if (need) { pos = 624; break; }
        }
#ifdef PG_HER_TIMING
        tLoop += clock64() - tw0;
#endif
        if (t == 0) { s_pos = pos; s_count = k; s_stop = need; }
      }
      __syncthreads();
#ifdef PG_HER_TIMING
      tWalk += clock64() - tc; tc = clock64();
#endif
      if (s_stop) { mt_twist(mt); if (t == 0) { s_pos = 0; s_tw_ok = 0; } __syncthreads(); }
#ifdef PG_HER_TIMING
      tTwist += clock64() - tc;
#endif
    }
    for (int k = t; k < cn; k += blockDim.x) {
      const int tt = s_cur[k] + (int)s_rng[k];
      H.s_t[c0 + k] = tt;
      H.s_trow[c0 + k] = (tt + s_st[k]) % rows;
    }
  }
  __syncthreads();
#ifdef PG_HER_TIMING
  if (t == 0) printf("her exact B=%d: phaseA %lld  stage %lld  walk(+temper) %lld  twist %lld  total %lld cycles; warp loop %lld cycles, %d rounds\n", B, tA1 - tA0, tStage, tWalk, tTwist, clock64() - tA0, tLoop, nRounds);
#endif
  if (t == 0) mt[624] = (uint32_t)s_pos;
  __syncthreads();
  for (int k = t; k < 625; k += blockDim.x) H.mt[k] = mt[k];
}

// Index sampling, device mode: counter-based hash draws (unbiased multiply-shift on 64 bits), fully parallel.
// NOT numpy-stream compatible; same distribution.
__global__ void k_her_indices_fast(HerParams H, int B, int nb_virtual, uint64_t seed, uint64_t counter) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= B) return;
  const int nv = H.n_valid[0], ne = H.n_envs, rows = H.rows;
  if (nv <= 0) { if (k == 0) H.status[0] = 1; return; }
  if (k == 0) H.status[0] = 0;
  uint64_t h = splitmix64(seed ^ splitmix64(counter * 0x9E3779B97F4A7C15ULL + (uint64_t)k));
  int idx = H.valid[(int)(((h >> 32) * (uint64_t)nv) >> 32)];
  int row = idx / ne, env = idx % ne;
  H.s_row[k] = row; H.s_env[k] = env;
  if (k < nb_virtual) {
    int st = H.ep_start[(size_t)row * ne + env], len = H.ep_length[(size_t)row * ne + env];
    int cur = ((row - st) % rows + rows) % rows;
    uint64_t h2 = splitmix64(h);
    const int tt = cur + (int)(((h2 >> 32) * (uint64_t)(len - cur)) >> 32);
    H.s_t[k] = tt;
    H.s_trow[k] = (tt + st) % rows;
  }
}

// ------------------------------------------------------------------------------------------ gather + relabel
// Gather + relabel.  The output row of a sample is the stored record itself (same field order, whole 64 B lines) with
// four fields patched: desired / next desired goal <- future achieved goal (virtual samples), reward <- recomputed,
// done <- done * (1 - timeout).  One warp moves kGatherPer records per trip as 16-byte vectors: all loads of the trip
// (records + future goals, located by the index kernels so nothing here depends on a loaded value) are issued before
// the first store, the records are patched in shared memory by spare lanes and leave as full-line float4 stores.
// HBM-bound: 1.15 KB read + written per sample (576 B record, a 32 B sector of the future record, 576 B out).
constexpr int kGatherWarps = 8, kGatherPer = 4;
__global__ void __launch_bounds__(kGatherWarps * 32) k_her_gather(HerParams H, Params P, HerBatch O, int B, int nb_virtual, int reward_env) {
  extern __shared__ float4 s_rec[];  // [warp][kGatherPer][rec_w / 4]
  if (H.status[0] != 0) return;      // nothing valid to sample from: leave the batch alone (the host raises)
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int ne = H.n_envs, w = H.rec_w, w4 = w >> 2;
  const int o_ag = her_off_ag(H), o_rew = her_off_rew(H);
  float4* mine = s_rec + (size_t)wib * kGatherPer * w4;
  const int stride = gridDim.x * kGatherWarps * kGatherPer;
  for (int base = (blockIdx.x * kGatherWarps + wib) * kGatherPer; base < B; base += stride) {
    float4 v0[kGatherPer], v1[kGatherPer];
    float fut[kGatherPer];
#pragma unroll
    for (int j = 0; j < kGatherPer; j++) {
      const int k = base + j;  // SB3 sample order: the first nb_virtual are virtual
      v0[j] = make_float4(0.f, 0.f, 0.f, 0.f); v1[j] = v0[j]; fut[j] = 0.f;
      if (k < B) {
        const int row = H.s_row[k], env = H.s_env[k];
        PG_CHECK(row >= 0 && row < H.rows && env >= 0 && env < ne && w4 <= 64);
        PG_CHECK(!(k < nb_virtual) || (H.s_trow[k] >= 0 && H.s_trow[k] < H.rows));
        const float4* __restrict__ rec = reinterpret_cast<const float4*>(H.rec + ((size_t)row * ne + env) * w);
        if (lane < w4) v0[j] = __ldg(rec + lane);
        if (lane + 32 < w4) v1[j] = __ldg(rec + lane + 32);
        if (k < nb_virtual && lane < 3)  // next achieved goal of the future transition of the same episode
          fut[j] = __ldg(H.rec + ((size_t)H.s_trow[k] * ne + env) * w + o_ag + 6 + lane);
      }
    }
#pragma unroll
    for (int j = 0; j < kGatherPer; j++) {
      if (lane < w4) mine[j * w4 + lane] = v0[j];
      if (lane + 32 < w4) mine[j * w4 + lane + 32] = v1[j];
    }
    __syncwarp();
#pragma unroll
    for (int j = 0; j < kGatherPer; j++) {
      if (base + j < nb_virtual && lane < 3) {
        float* r = reinterpret_cast<float*>(mine + j * w4);
        r[o_ag + 3 + lane] = fut[j]; r[o_ag + 9 + lane] = fut[j];
      }
    }
    __syncwarp();
    if (lane < kGatherPer && base + lane < B) {
      float* r = reinterpret_cast<float*>(mine + lane * w4);
      if (base + lane < nb_virtual) {
        // virtual reward = env.compute_reward(next_achieved_goal, new_goal) of environment `reward_env` (SB3: indices=[0])
        const float nag[3] = {r[o_ag + 6], r[o_ag + 7], r[o_ag + 8]}, g[3] = {r[o_ag + 3], r[o_ag + 4], r[o_ag + 5]};
        r[o_rew] = reward_fn(P, reward_env, nag, g);
      }
      r[o_rew + 1] = r[o_rew + 1] * (1.f - r[o_rew + 2]);  // dones * (1 - timeouts)
    }
    __syncwarp();
#pragma unroll
    for (int j = 0; j < kGatherPer; j++) {
      const int k = base + j;
      if (k < B) {
        const int out = k < nb_virtual ? (B - nb_virtual) + k : k - nb_virtual;  // torch.cat((real, virtual))
        PG_CHECK(out >= 0 && out < B);
        float4* dst = reinterpret_cast<float4*>(O.packed + (size_t)out * w);
        if (lane < w4) dst[lane] = mine[j * w4 + lane];
        if (lane + 32 < w4) dst[lane + 32] = mine[j * w4 + lane + 32];
      }
    }
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------------------ launchers
void launch_her_add(const HerParams& H, const float* obs, const float* ag, const float* dg, const float* nobs, const float* nag,
                    const float* ndg, const float* act, const float* rew, const uint8_t* done, const uint8_t* timeout, cudaStream_t s) {
  k_her_add_book<<<(H.n_envs + 127) / 128, 128, 0, s>>>(H, done);
  int64_t tot = (int64_t)H.n_envs * (2 * H.obs_dim + 12 + H.act_dim + 3);
  k_her_add_data<<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(H, obs, ag, dg, nobs, nag, ndg, act, rew, done, timeout);
  k_her_advance<<<1, 1, 0, s>>>(H);
}
void launch_her_compact(const HerParams& H, int total, cudaStream_t s) {
  const int nb = (total + kHerBlock - 1) / kHerBlock;
  k_her_count<<<nb, kHerBlock, 0, s>>>(H, total);
  k_her_scan_blocks<<<1, 1024, 0, s>>>(H, nb);
  k_her_scatter<<<nb, kHerBlock, 0, s>>>(H, total);
}
void launch_her_indices(const HerParams& H, int B, int nbv, bool exact, uint64_t seed, uint64_t counter, cudaStream_t s) {
  if (exact) k_her_indices_exact<<<1, 256, 0, s>>>(H, B, nbv);
  else k_her_indices_fast<<<(B + 255) / 256, 256, 0, s>>>(H, B, nbv, seed, counter);
}
void launch_her_gather(const HerParams& H, const Params& P, const HerBatch& O, int B, int nbv, int reward_env, int sm_count, cudaStream_t s) {
  const int per_block = kGatherWarps * kGatherPer;
  int blocks = (B + per_block - 1) / per_block, cap = sm_count * 8;  // persistent: 8 blocks (64 warps) per SM
  if (blocks > cap) blocks = cap;
  const size_t smem = (size_t)per_block * H.rec_w * sizeof(float);  // 18 KB at 576 B records; rec_w <= 256 floats checked at create
  k_her_gather<<<blocks, kGatherWarps * 32, smem, s>>>(H, P, O, B, nbv, reward_env);
}

}  // namespace pg
