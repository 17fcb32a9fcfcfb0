// pg_tu_obj.cu -- the Push / PickAndPlace step kernel (arm + free box + frictional contacts, pg_contact.cuh)
#include "pg_kernels.cuh"
#include "pg_launch.h"

namespace pg {

// Launch shape: BASELINE configs[3] runs 16 384 environments per GPU = 111 per SM.  One environment per thread, ONE block of
// 112 threads per SM (147 blocks): a thread may then use up to 255 registers (the dynamics + solver state stays in
// registers) and the block owns the SM's shared memory, which holds what used to spill: the 12-slot contact store
// (23 / 41 floats per slot) and the task-space matrix X -- 462 floats per thread, 207 KB per block, structure-of-arrays over the block.
constexpr int kObjBlock = 112;
constexpr int kObjSmemFloats = (kCStoreFloats + kXG) * kObjBlock;
__global__ void __launch_bounds__(kObjBlock, 1) k_step_obj(const Params P, const float* __restrict__ actions) {
  extern __shared__ float s_obj[];
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = e < P.n;  // no block barriers here: with one warp per scheduler the warps run best unsynchronised
  CStore S; S.base = s_obj + threadIdx.x; S.stride = kObjBlock;
  float* XG = s_obj + kCStoreFloats * kObjBlock + threadIdx.x;
  float q[7], qd[7], cq[7], sq[7], tgt[7];
  bool velocity_mode = false, unit_imp = false;
  ObjState B;
  if (valid) {
    step_begin(P, actions, e, q, qd, tgt, 1, velocity_mode, unit_imp);
#pragma unroll
    for (int d = 0; d < 7; d++) sincosf(q[d], &sq[d], &cq[d]);
    B = load_obj(P, e);
  }
  const V3 base = v3(P.base[0], P.base[1], P.base[2]);
  for (int k = 0; k < P.n_substeps; k++)
    world_substep<false>(valid, velocity_mode, unit_imp, base, P.world, q, qd, cq, sq, tgt, B, S, XG, kObjBlock);
  if (!valid) return;
  store_obj(P, e, B);
  step_end(P, e, q, qd, false);
}

void launch_step_obj(const Params& P, const float* actions, cudaStream_t s) {
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  const int smem = kObjSmemFloats * (int)sizeof(float);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {  // opt in to > 48 KB dynamic shared memory, once per device
    cudaFuncSetAttribute(k_step_obj, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    attr_set[dev] = true;
  }
  k_step_obj<<<(P.n + kObjBlock - 1) / kObjBlock, kObjBlock, smem, s>>>(P, actions);
}
}  // namespace pg
