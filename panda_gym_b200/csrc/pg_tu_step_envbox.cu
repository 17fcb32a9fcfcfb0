// pg_tu_step_envbox.cu -- the step kernel with per-environment cuboid sizes (scenario reachao_rand_shape, reach_ao.py:606-608,
// 1084-1089): same body as pg_tu_step.cu, its collision / observation passes read each environment's own box shapes.
#define PG_ENVBOX 1
#include "pg_kernels.cuh"
#include "pg_launch.h"

namespace pg {

#define PG_STEP_KERNEL_NAME k_step_envbox
#include "pg_step_body.cuh"

void launch_step_envbox(const Params& P, const float* actions, cudaStream_t s) {
  static bool attr_set[64] = {};  // > 48 KB of dynamic shared memory needs the opt-in, once per device
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    cudaFuncSetAttribute(k_step_envbox, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    attr_set[dev] = true;
  }
  k_step_envbox<<<(P.n + kStepBlock - 1) / kStepBlock, kStepBlock, kStepSmem, s>>>(P, actions);
}
}  // namespace pg
