// pg_her.cuh -- on-device Hindsight-Experience-Replay buffer: storage, episode bookkeeping, index sampling,
// future-goal relabelling + reward recompute, batch gather.
//
// Replaces stable_baselines3.her.her_replay_buffer.HerReplayBuffer (SB3 >= 2.0 semantics; the reference uses it
// through classes/train_config.py:2,15 with default kwargs: n_sampled_goal = 4, goal_selection_strategy = "future").
// Contract: SURVEY.md App. D.  Index sampling in "numpy" mode consumes a legacy MT19937 stream exactly like
// np.random.choice(valid, B) followed by np.random.randint(cur[], ep_len[]) (32-bit masked rejection), so that
// (row, env, t) triples are bit-identical to SB3's for the same np.random state.
#pragma once
#include <stdint.h>
#include "pg_kernels.cuh"

namespace pg {

struct HerParams {
  int rows, n_envs, obs_dim, act_dim;
  // storage: ONE record per transition, [rows][n_envs][rec_w] f32, rec_w = used width rounded up to 16 floats (64 B):
  //   observation od | next observation od | achieved 3 | desired 3 | next achieved 3 | next desired 3 | action ad |
  //   reward | done | timeout | ep_start (int bits) | pad
  // A sampled transition is one contiguous, 64 B-aligned read (576 B at obs 56 / act 7) instead of ten scattered
  // fields of ten arrays (each a separate 32 B sector: measured 2x read amplification).
  float* rec;
  int rec_w;
  int *ep_start, *ep_length;  // [rows][n_envs] (bookkeeping + valid-index compaction; ep_start is mirrored into the record)
  int* cur_ep_start;          // [n_envs]
  int* cursor;                // [0] pos, [1] full
  // sampling scratch
  int* valid;                 // compacted flat indices (row * n_envs + env) of transitions with ep_length > 0
  int* n_valid;               // [0]
  int* block_counts;          // per-block counts for the compaction
  uint32_t* mt;               // [624] MT19937 state, [624] = pos
  uint32_t* raw;              // raw tempered words for one sample() call
  int raw_cap;
  int *s_row, *s_env, *s_t;   // last sampled indices (in SB3's order: virtual first, then real)
  int* s_trow;                // storage row of the future transition of each virtual sample ((t + ep_start) % rows)
  int* status;                // [0] != 0 when sample() could not run (no valid transition)
};

constexpr int kHerBlock = 256;  // block size of the valid-index compaction

// record field offsets (floats)
__host__ __device__ __forceinline__ int her_used(int od, int ad) { return 2 * od + 12 + ad + 4; }
__device__ __forceinline__ int her_off_ag(const HerParams& H) { return 2 * H.obs_dim; }
__device__ __forceinline__ int her_off_act(const HerParams& H) { return 2 * H.obs_dim + 12; }
__device__ __forceinline__ int her_off_rew(const HerParams& H) { return 2 * H.obs_dim + 12 + H.act_dim; }
__device__ __forceinline__ int her_off_eps(const HerParams& H) { return 2 * H.obs_dim + 12 + H.act_dim + 3; }

struct HerBatch {  // output: [batch][rec_w] rows in the record's field order, order = torch.cat((real, virtual)) like SB3
  float* packed;
};

}  // namespace pg
