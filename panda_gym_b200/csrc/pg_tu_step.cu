// pg_tu_step.cu -- the ReachAO / Reach step kernel (RobotTaskEnv.step, core.py:352-368) and its raw-substep parity twin
#define PG_ENVBOX 0  // this unit's flag pass never looks for per-environment cuboid sizes (pg_scene.cuh); the twin is pg_tu_step_envbox.cu
#include "pg_kernels.cuh"
#include "pg_launch.h"

namespace pg {

#define PG_STEP_KERNEL_NAME k_step
#include "pg_step_body.cuh"

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
  if (P.scene.env_box) { launch_step_envbox(P, actions, s); return; }  // reachao_rand_shape
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

#ifdef PG_SWEEP_PROBE
// variant build only (tools/sweep_probe.py): read and clear the per-substep, per-environment sweep-pair counters of k_step
// out_host: int [20][148 * 512]
extern "C" int pg_debug_sweeps_read(int* out_host) {
  if (cudaMemcpyFromSymbol(out_host, pg::g_sweeps, sizeof(int) * 20 * 148 * 512) != cudaSuccess) return -1;
  static int zeros[20 * 148 * 512];
  return (int)cudaMemcpyToSymbol(pg::g_sweeps, zeros, sizeof zeros);
}
#endif
