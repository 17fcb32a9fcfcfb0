// pg_tu_step.cu -- the ReachAO / Reach step kernel (RobotTaskEnv.step, core.py:352-368) and its raw-substep parity twin
#include "pg_kernels.cuh"
#include "pg_launch.h"

namespace pg {

// One environment per thread, one 448-thread block per SM.
__global__ void PG_STEP_BOUNDS k_step(const Params P, const float* __restrict__ actions) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = P.n;
  const bool valid = e < n;  // every thread of the block stays for the barriers of the substep loop
  // joint targets and angles are touched twice per substep: kept out of the register file; q0 / slack / ctr0 = CullState
  extern __shared__ float s_step[];
  float* tgt = s_step + threadIdx.x;
  float* qs = s_step + 7 * kStepBlock + threadIdx.x;
  float* q0 = s_step + 14 * kStepBlock + threadIdx.x;
  float* slack = s_step + 21 * kStepBlock + threadIdx.x;
  float* ctr0 = s_step + 30 * kStepBlock + threadIdx.x;
  float q[7], qd[7], cq[7], sq[7];
  bool velocity_mode = false, unit_imp = false;
  const bool cull = P.task == 1 && P.scene.skip_far != 0;
  if (valid) {
    step_begin(P, actions, e, q, qd, tgt, kStepBlock, velocity_mode, unit_imp);
#pragma unroll
    for (int d = 0; d < 7; d++) { sincosf(q[d], &sq[d], &cq[d]); qs[d * kStepBlock] = q[d]; q0[d * kStepBlock] = q[d]; }
    if (cull) {
      bool same = true;  // d_link belongs to this very pose (not the case after a state write without pg_observe)
#pragma unroll
      for (int d = 0; d < 7; d++) same &= P.q_dl[(size_t)d * n + e] == q[d];
#pragma unroll
      for (int i = 0; i < 9; i++) slack[i * kStepBlock] = same ? P.d_link[(size_t)i * n + e] - kSkipMargin : -1.f;
      scene_centres(v3(P.base[0], P.base[1], P.base[2]), cq, sq, ctr0, kStepBlock);
    }
  }
  // --- sim.step(): n_substeps x stepSimulation, collision check after each (ReachAO); a collided environment
  //     stops simulating (reach_ao.py:186-188) but its thread keeps meeting the block barriers
  bool collided = false, active = valid;
  const V3 base = v3(P.base[0], P.base[1], P.base[2]);
  CullState CS; CS.q = qs; CS.q0 = q0; CS.slack = slack; CS.ctr0 = ctr0; CS.stride = kStepBlock;
  for (int k = 0; k < P.n_substeps; k++) {
    substep<kStepSync>(active, velocity_mode, unit_imp, qs, qd, cq, sq, tgt, kStepBlock);
    if (P.task == 1) {
      bool hit = scene_pass<false, kStepSync>(active, P.scene, P.obs_pos_f, n, e, base, cq, sq, nullptr, nullptr, cull ? &CS : nullptr);
      if (active && hit) { collided = true; active = false; }
    }
  }
  if (!valid) return;
#pragma unroll
  for (int d = 0; d < 7; d++) q[d] = qs[d * kStepBlock];
  step_end(P, e, q, qd, collided);
}

// ------------------------------------------------------------------------------------------------ debug / parity kernels
__global__ void PG_STEP_BOUNDS k_debug_substeps(int n, float* q_io, float* qd_io, const float* tgt_in, float f0, float f1, float f2,
                                                           float f3, float f4, float f5, float f6, int nsub, int velocity) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = e < n;
  extern __shared__ float s_step[];
  float* t = s_step + threadIdx.x;
  float* q = s_step + 7 * kStepBlock + threadIdx.x;  // element d at q[d * kStepBlock]
  float qd[7], c[7], s[7];
  if (valid) {
#pragma unroll
    for (int d = 0; d < 7; d++) { q[d * kStepBlock] = q_io[(size_t)d * n + e]; qd[d] = qd_io[(size_t)d * n + e]; t[d * kStepBlock] = tgt_in[(size_t)d * n + e]; sincosf(q[d * kStepBlock], &s[d], &c[d]); }
  }
  for (int k = 0; k < nsub; k++) {
    // parity kernel: re-evaluate cos/sin exactly every 20 substeps like the env step does
    if (valid && k > 0 && (k % 20) == 0) {
#pragma unroll
      for (int d = 0; d < 7; d++) sincosf(q[d * kStepBlock], &s[d], &c[d]);
    }
    substep<kStepSync>(valid, velocity != 0, false, q, qd, c, s, t, kStepBlock);
  }
  if (!valid) return;
#pragma unroll
  for (int d = 0; d < 7; d++) { q_io[(size_t)d * n + e] = q[d * kStepBlock]; qd_io[(size_t)d * n + e] = qd[d]; }
}

constexpr int kStepSmem = 57 * kStepBlock * (int)sizeof(float);  // tgt 7 | q 7 | q0 7 | slack 9 | ctr0 27
static void step_smem_opt_in() {  // > 48 KB of dynamic shared memory needs the opt-in, once per device
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    cudaFuncSetAttribute(k_step, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    cudaFuncSetAttribute(k_debug_substeps, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    attr_set[dev] = true;
  }
}
void launch_step(const Params& P, const float* actions, cudaStream_t s) {
  step_smem_opt_in();
  k_step<<<(P.n + kStepBlock - 1) / kStepBlock, kStepBlock, kStepSmem, s>>>(P, actions);
}
void launch_debug_substeps(int n, float* q, float* qd, const float* tgt, const float* f, int n_sub, int velocity_mode, cudaStream_t s) {
  step_smem_opt_in();
  k_debug_substeps<<<(n + kStepBlock - 1) / kStepBlock, kStepBlock, kStepSmem, s>>>(n, q, qd, tgt, f[0], f[1], f[2], f[3], f[4], f[5], f[6], n_sub, velocity_mode);
}
cudaError_t step_kernel_attrs(int* out) {
  cudaFuncAttributes a;
  int occ = 0;
  cudaError_t e = cudaFuncGetAttributes(&a, k_step);
  if (e != cudaSuccess) return e;
  step_smem_opt_in();
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_step, kStepBlock, kStepSmem);
  out[0] = a.numRegs; out[1] = a.maxThreadsPerBlock; out[2] = (int)a.sharedSizeBytes; out[3] = (int)a.localSizeBytes;
  out[4] = kStepBlock; out[5] = occ;
  return cudaSuccess;
}
}  // namespace pg
