// pg_tu_her.cu -- kernels of the on-device HER replay buffer (contract: pg_her.cuh)
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

// Index sampling, numpy-exact.  ONE block of 256 threads (blockDim >= 128 so each twist phase fits 3 values/thread).
//   phase A: np.random.choice(valid, B) == randint(0, n_valid, B): masked rejection on 32-bit draws, all elements
//            share one mask -> accept flags + ordered compaction, chunk by chunk;
//   phase B: np.random.randint(cur[], ep_len[]) for the first nb_virtual samples: per-element mask, elements in
//            order, an element with range 0 consumes no draw -> a sequential automaton over the raw stream.
// Raw words are produced on demand; the MT state left behind equals numpy's after the same two calls.
__global__ void __launch_bounds__(256) k_her_indices_exact(HerParams H, int B, int nb_virtual) {
  __shared__ uint32_t mt[625];
  __shared__ int s_scan[256];
  __shared__ int s_count, s_stop, s_pos;
  const int t = threadIdx.x;
  for (int k = t; k < 625; k += blockDim.x) mt[k] = H.mt[k];
  __syncthreads();
  const int nv = H.n_valid[0];
  if (nv <= 0) { if (t == 0) H.status[0] = 1; return; }
  if (t == 0) { H.status[0] = 0; s_count = 0; s_pos = (int)mt[624]; }
  __syncthreads();
  const int rows = H.rows, ne = H.n_envs;
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
      s_scan[t] = f;
      __syncthreads();
      for (int o = 1; o < 256; o <<= 1) { int x = t >= o ? s_scan[t - o] : 0; __syncthreads(); s_scan[t] += x; __syncthreads(); }
      int before = s_count, mine = before + s_scan[t] - 1;  // index of this accepted draw
      if (t == 0) s_stop = -1;
      __syncthreads();
      if (f && mine < B) { int idx = H.valid[w]; H.s_row[mine] = idx / ne; H.s_env[mine] = idx % ne; if (mine == B - 1) s_stop = t; }
      __syncthreads();
      if (t == 0) {
        if (s_stop >= 0) { s_pos = pos + s_stop + 1; s_count = B; }
        else { s_pos = pos + chunk; s_count = before + s_scan[255]; }
      }
      __syncthreads();
      if (s_count >= B) break;
    }
  }
  // ---- phase B (future strategy): t = randint(cur, ep_len) per virtual sample, sequential consumption
  __syncthreads();
  int done_el = 0;
  while (done_el < nb_virtual) {
    if (s_pos >= 624) { __syncthreads(); mt_twist(mt); if (t == 0) s_pos = 0; __syncthreads(); }
    if (t == 0) {
      int pos = s_pos, k = done_el;
      while (k < nb_virtual) {
        int row = H.s_row[k], env = H.s_env[k];
        int st = H.ep_start[(size_t)row * ne + env], len = H.ep_length[(size_t)row * ne + env];
        int cur = ((row - st) % rows + rows) % rows;
        uint32_t r = (uint32_t)(len - 1 - cur);  // high - 1 - low
        uint32_t v = 0;
        if (r != 0) {
          uint32_t m = r; m |= m >> 1; m |= m >> 2; m |= m >> 4; m |= m >> 8; m |= m >> 16;
          bool got = false;
          while (pos < 624) { v = mt_temper(mt[pos++]) & m; if (v <= r) { got = true; break; } }
          if (!got) break;  // state exhausted mid-element: twist and retry this element
        }
        H.s_t[k] = cur + (int)v;
        k++;
      }
      s_pos = pos; s_count = k;
    }
    __syncthreads();
    done_el = s_count;
    __syncthreads();
  }
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
    H.s_t[k] = cur + (int)(((h2 >> 32) * (uint64_t)(len - cur)) >> 32);
  }
}

// ------------------------------------------------------------------------------------------ gather + relabel
// one warp per sampled transition
__global__ void k_her_gather(HerParams H, Params P, HerBatch O, int B, int nb_virtual, int reward_env) {
  int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  const int k = warp;                       // SB3 sample order: first nb_virtual are virtual
  const bool is_virtual = k < nb_virtual;
  const int out = is_virtual ? (B - nb_virtual) + k : k - nb_virtual;
  const int ne = H.n_envs, od = H.obs_dim, ad = H.act_dim, rows = H.rows, w = H.rec_w;
  const int row = H.s_row[k], env = H.s_env[k];
  const float* __restrict__ rec = H.rec + ((size_t)row * ne + env) * w;
  const int o_ag = her_off_ag(H), o_act = her_off_act(H), o_rew = her_off_rew(H);
  // field-wise copy out of the contiguous record (warp-uniform loops: no divergence; the whole record is 9 x 64 B lines)
  for (int i = lane; i < od; i += 32) {
    const float a = rec[i], b = rec[od + i];
    O.obs[(size_t)out * od + i] = a; O.next_obs[(size_t)out * od + i] = b;
  }
  for (int i = lane; i < ad; i += 32) O.action[(size_t)out * ad + i] = rec[o_act + i];
  // new goal: the achieved goal of a later transition of the same episode (virtual) or the stored desired goal
  float g[3], nag[3];
  if (is_virtual) {
    const int st = __float_as_int(rec[o_rew + 3]);
    const int trow = (H.s_t[k] + st) % rows;
    const float* fut = H.rec + ((size_t)trow * ne + env) * w + o_ag + 6;  // its next achieved goal
    g[0] = fut[0]; g[1] = fut[1]; g[2] = fut[2];
  } else { g[0] = rec[o_ag + 3]; g[1] = rec[o_ag + 4]; g[2] = rec[o_ag + 5]; }
  nag[0] = rec[o_ag + 6]; nag[1] = rec[o_ag + 7]; nag[2] = rec[o_ag + 8];
  if (lane < 3) {
    O.ag[(size_t)out * 3 + lane] = rec[o_ag + lane];
    O.next_ag[(size_t)out * 3 + lane] = nag[lane];
    O.dg[(size_t)out * 3 + lane] = g[lane];
    O.next_dg[(size_t)out * 3 + lane] = is_virtual ? g[lane] : rec[o_ag + 9 + lane];
  }
  if (lane == 0) {
    // virtual reward = env.compute_reward(next_achieved_goal, new_goal) of environment `reward_env` (SB3: indices=[0])
    O.reward[out] = is_virtual ? reward_fn(P, reward_env, nag, g) : rec[o_rew];
    O.done[out] = rec[o_rew + 1] * (1.f - rec[o_rew + 2]);  // dones * (1 - timeouts)
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
void launch_her_gather(const HerParams& H, const Params& P, const HerBatch& O, int B, int nbv, int reward_env, cudaStream_t s) {
  k_her_gather<<<(unsigned)(((int64_t)B * 32 + 255) / 256), 256, 0, s>>>(H, P, O, B, nbv, reward_env);
}

}  // namespace pg
