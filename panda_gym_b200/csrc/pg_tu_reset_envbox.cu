// pg_tu_reset_envbox.cu -- reset / observation-refresh kernels with per-environment cuboid sizes (scenario reachao_rand_shape,
// reach_ao.py:606-608, 969-979, 1084-1089): same bodies as pg_tu_reset.cu, the scene pass reads each environment's own box shapes.
#define PG_ENVBOX 1
#include "pg_reset.cuh"
#include "pg_launch.h"

namespace pg {

#define PG_RESET_KERNEL k_reset_envbox
#define PG_OBSERVE_KERNEL k_observe_envbox
#include "pg_reset_body.cuh"

void launch_reset_envbox(const Params& P, bool list, const uint8_t* mask, const uint64_t* seeds, uint64_t first_env, int sm_count, cudaStream_t s) {
  const bool gp = P.task == 1 && P.st.pose_active && !P.st.random_base;
  if (list) {
    if (gp) k_reset_envbox<true, true><<<sm_count * 8, kBlock, 0, s>>>(P, nullptr, nullptr, first_env);
    else k_reset_envbox<true, false><<<sm_count * 8, kBlock, 0, s>>>(P, nullptr, nullptr, first_env);
  } else {
    if (gp) k_reset_envbox<false, true><<<(P.n + kBlock - 1) / kBlock, kBlock, 0, s>>>(P, mask, seeds, first_env);
    else k_reset_envbox<false, false><<<(P.n + kBlock - 1) / kBlock, kBlock, 0, s>>>(P, mask, seeds, first_env);
  }
}
void launch_observe_envbox(const Params& P, cudaStream_t s) { k_observe_envbox<<<(P.n + kBlock - 1) / kBlock, kBlock, 0, s>>>(P); }
}  // namespace pg
