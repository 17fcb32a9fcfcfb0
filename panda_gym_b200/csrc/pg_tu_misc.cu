// pg_tu_misc.cu -- small kernels: obstacle mirror, vectorised compute_reward, synthetic actions, evaluation metrics, manipulability, parity probes
#include "pg_kernels.cuh"
#include "pg_launch.h"

namespace pg {

__global__ void k_mirror_obstacles(const Params P) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (int64_t)kMaxObst * 3 * P.n) P.obs_pos_f[i] = (float)P.obs_pos[i];
  if (P.st.random_size && i < (int64_t)kMaxObst * P.n) {  // fp32 mirror of the per-environment cuboid sizes (pg_reset.cuh: store_env_box)
    const int j = (int)(i / P.n), e = (int)(i % P.n);
    const double sx = P.env_size[(size_t)(j * 3) * P.n + e], sy = P.env_size[(size_t)(j * 3 + 1) * P.n + e], sz = P.env_size[(size_t)(j * 3 + 2) * P.n + e];
    float* b = P.env_box + (size_t)(j * 5) * P.n + e;
    b[0] = (float)sx; b[P.n] = (float)sy; b[(size_t)2 * P.n] = (float)sz;
    b[(size_t)3 * P.n] = (float)fmin(0.04, 0.1 * fmin(sx, fmin(sy, sz)));
    b[(size_t)4 * P.n] = sqrtf((float)sx * (float)sx + (float)sy * (float)sy + (float)sz * (float)sz);
  }
}

__global__ void k_reward(const Params P, const float* __restrict__ ag, const float* __restrict__ dg, float* __restrict__ out, int64_t m, int src) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  float a[3] = {ag[3 * i], ag[3 * i + 1], ag[3 * i + 2]}, g[3] = {dg[3 * i], dg[3 * i + 1], dg[3 * i + 2]};
  out[i] = reward_fn(P, src, a, g);
}

__global__ void k_random_actions(float* __restrict__ out, int64_t m, uint64_t seed, uint64_t counter) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  uint64_t h = splitmix64(seed ^ splitmix64(counter * 0x100000001B3ULL + (uint64_t)i));
  out[i] = (float)((double)(h >> 11) * (2.0 / 9007199254740992.0) - 1.0);
}

__global__ void k_debug_minv(int n, const float* q_in, float* out) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  float c[7], s[7], M[28];
#pragma unroll
  for (int d = 0; d < 7; d++) sincosf(q_in[(size_t)d * n + e], &s[d], &c[d]);
  crba(c, s, M);
  spd_inverse7(M);
#pragma unroll
  for (int k = 0; k < 28; k++) out[(size_t)k * n + e] = M[k];
}
// The per-step bookkeeping of evaluation/evaluate.py:perform_benchmark (:236-262) and its episode-end accounting
// (:264-283), for every environment, on the state k_step left behind and BEFORE the auto-reset overwrites it:
//   per step     total_effort += ||current_joint_acceleration||   (reach_ao.py:1299-1302)
//                jerk.append(||current_joint_jerk||)              (:1304-1305)
//                total_manipulability += task.manipulability      (:1311, :1381 -> panda.py:314-317 at the new q)
//                ee_speed.append(||get_ee_velocity()||)           (observation[3:6])
//   episode end  is_success -> success bucket (length, effort / length, manipulability / length, all jerk and speed
//                samples); else is_truncated -> collision; else timeout; reward and length always
__global__ void __launch_bounds__(128) k_eval_metrics(const Params P) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x, n = P.n;
  if (e >= n) return;
  float c[7], s[7], a2 = 0.f, j2 = 0.f;
#pragma unroll
  for (int d = 0; d < 7; d++) {
    sincosf(P.q[(size_t)d * n + e], &s[d], &c[d]);
    float a = P.cur_acc[(size_t)d * n + e], j = P.jerk[(size_t)d * n + e];
    a2 = fmaf(a, a, a2); j2 = fmaf(j, j, j2);
  }
  const float* o = P.obs + (size_t)e * P.obs_dim;
  const float speed = sqrtf(o[3] * o[3] + o[4] * o[4] + o[5] * o[5]);
  float effort = P.ev_acc[e] + sqrtf(a2), manip = P.ev_acc[(size_t)n + e] + manipulability_cs(c, s);
  float jerk = P.ev_acc[(size_t)2 * n + e] + sqrtf(j2), sp = P.ev_acc[(size_t)3 * n + e] + speed;
  int len = P.ev_len[e] + 1;
  float ret = P.ev_ret[e] + P.reward[e];
  if (P.done[e]) {
    const uint8_t f = P.flags[e];
    atomicAdd(&P.ev[6], 1.0); atomicAdd(&P.ev[7], (double)len); atomicAdd(&P.ev[9], (double)ret);
    if (f & 1) {
      atomicAdd(&P.ev[0], 1.0); atomicAdd(&P.ev[1], (double)len);
      atomicAdd(&P.ev[2], (double)effort / len); atomicAdd(&P.ev[3], (double)manip / len);
      atomicAdd(&P.ev[4], (double)jerk); atomicAdd(&P.ev[5], (double)sp);
    } else if (f & 4) atomicAdd(&P.ev[8], 1.0);
    else atomicAdd(&P.ev[10], 1.0);
    effort = manip = jerk = sp = ret = 0.f; len = 0;
  }
  P.ev_acc[e] = effort; P.ev_acc[(size_t)n + e] = manip; P.ev_acc[(size_t)2 * n + e] = jerk; P.ev_acc[(size_t)3 * n + e] = sp;
  P.ev_len[e] = len; P.ev_ret[e] = ret;
}

// Panda.get_manipulability (panda.py:314-317): roboticstoolbox Panda().manipulability(q, axes="trans") = Yoshikawa index
// sqrt(det(Jv Jv^T)) of the 3 x 7 translational Jacobian at the toolbox model's end-effector link panda_link8 (flange =
// link7 frame + (0, 0, 0.107)).  compute_reward refreshes it every step (reach_ao.py:1311, :1381); it feeds no reward.
// q f32 [7][n] -> out f32 [n].  Same formula as oracle pgo_arm_manipulability.
__global__ void k_manipulability(int n, const float* __restrict__ q_in, float* __restrict__ out) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  float c[7], s[7];
#pragma unroll
  for (int d = 0; d < 7; d++) sincosf(q_in[(size_t)d * n + e], &s[d], &c[d]);
  ArmKin K;
  arm_kin(v3(0.f, 0.f, 0.f), c, s, K);
  const V3 p = frame_point(K.F7, 0.f, 0.f, 0.107f);
  // det(Jv Jv^T) by Cauchy-Binet = sum over the 35 column triples of det(J_S)^2: no cancellation between the terms, so the
  // index keeps its relative accuracy in fp32 near singular poses (the 3 x 3 Gram determinant does not)
  V3 J[7];
#pragma unroll
  for (int d = 0; d < 7; d++) J[d] = cross(K.ax[d], p - K.org[d]);
  float det = 0.f;
#pragma unroll
  for (int i = 0; i < 5; i++)
#pragma unroll
    for (int j = i + 1; j < 6; j++) {
      V3 cij = cross(J[i], J[j]);
#pragma unroll
      for (int k = j + 1; k < 7; k++) { float m = dot(cij, J[k]); det = fmaf(m, m, det); }
    }
  out[e] = det > 0.f ? sqrtf(det) : 0.f;
}
__global__ void k_debug_ik(int n, const float* q_in, const float* tgt, float bx, float by, float bz, float* out) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  float q[7], o[7];
#pragma unroll
  for (int d = 0; d < 7; d++) q[d] = q_in[(size_t)d * n + e];
  ik_dls(v3(bx, by, bz), q, v3(tgt[e], tgt[(size_t)n + e], tgt[(size_t)2 * n + e]), o);
#pragma unroll
  for (int d = 0; d < 7; d++) out[(size_t)d * n + e] = o[d];
}

__global__ void __launch_bounds__(256) k_fma_peak(float* out, int iters, float a, float b) {
  float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
    x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
    x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
  }
  float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  if (s == 123.456f) out[0] = s;
}

void launch_mirror_obstacles(const Params& P, cudaStream_t s) {
  k_mirror_obstacles<<<(unsigned)(((size_t)kMaxObst * 3 * P.n + 255) / 256), 256, 0, s>>>(P);
}
void launch_reward(const Params& P, const float* ag, const float* dg, float* out, int64_t n, int src, cudaStream_t s) {
  k_reward<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(P, ag, dg, out, n, src);
}
void launch_random_actions(float* out, int64_t m, uint64_t seed, uint64_t counter, cudaStream_t s) {
  k_random_actions<<<(unsigned)((m + 255) / 256), 256, 0, s>>>(out, m, seed, counter);
}
void launch_eval_metrics(const Params& P, cudaStream_t s) { k_eval_metrics<<<(P.n + 127) / 128, 128, 0, s>>>(P); }
void launch_manipulability(int n, const float* q, float* out, cudaStream_t s) { k_manipulability<<<(n + 127) / 128, 128, 0, s>>>(n, q, out); }
void launch_debug_minv(int n, const float* q, float* out, cudaStream_t s) { k_debug_minv<<<(n + 127) / 128, 128, 0, s>>>(n, q, out); }
void launch_debug_ik(int n, const float* q, const float* tgt, float bx, float by, float bz, float* out, cudaStream_t s) {
  k_debug_ik<<<(n + 127) / 128, 128, 0, s>>>(n, q, tgt, bx, by, bz, out);
}
void launch_fma_peak(int blocks, float* out, int iters, cudaStream_t s) { k_fma_peak<<<blocks, 256, 0, s>>>(out, iters, 0.999f, 0.001f); }
}  // namespace pg
