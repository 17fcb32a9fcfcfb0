// pg_tu_obj.cu -- the Push / PickAndPlace step kernel (arm + free box + frictional contacts, pg_contact.cuh)
#include "pg_kernels.cuh"
#include "pg_launch.h"

namespace pg {

// Launch shape: BASELINE configs[3] runs 16 384 environments per GPU = 512 warps for 592 warp schedulers, so the kernel is
// latency-bound whatever the block size; 32-thread blocks at least spread the warps over all 148 SMs (128-thread blocks
// left 20 SMs idle) and give each warp a whole L1 for its solver rows.
constexpr int kObjBlock = 32;
__global__ void __launch_bounds__(kObjBlock) k_step_obj(const Params P, const float* __restrict__ actions) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= P.n) return;
  float q[7], qd[7], cq[7], sq[7], tgt[7];
  bool velocity_mode = false, unit_imp = false;
  step_begin(P, actions, e, q, qd, tgt, 1, velocity_mode, unit_imp);
#pragma unroll
  for (int d = 0; d < 7; d++) sincosf(q[d], &sq[d], &cq[d]);
  ObjState B = load_obj(P, e);
  const V3 base = v3(P.base[0], P.base[1], P.base[2]);
  for (int k = 0; k < P.n_substeps; k++) world_substep(velocity_mode, unit_imp, base, P.world, q, qd, cq, sq, tgt, B);
  store_obj(P, e, B);
  step_end(P, e, q, qd, false);
}

void launch_step_obj(const Params& P, const float* actions, cudaStream_t s) {
  k_step_obj<<<(P.n + kObjBlock - 1) / kObjBlock, kObjBlock, 0, s>>>(P, actions);
}
}  // namespace pg
