// pg_tu_reset.cu -- dispatch to the reset / observation-refresh kernel variants (pg_tu_reset_e?g?.cu) and the reset-pose table
#include "pg_reset.cuh"
#include "pg_launch.h"

namespace pg {

void launch_reset_all(const Params& P, const uint8_t* mask, const uint64_t* seeds, uint64_t first_env, cudaStream_t s) {
  const bool g = P.task == 1 && P.st.pose_active && !P.st.random_base;  // a start-pose randomizer that puts the arm anywhere
  (P.scene.env_box ? (g ? launch_reset_e1g1 : launch_reset_e1g0) : (g ? launch_reset_e0g1 : launch_reset_e0g0))(P, false, mask, seeds, first_env, 0, s);
}
void launch_reset_list(const Params& P, uint64_t first_env, int sm_count, cudaStream_t s) {
  const bool g = P.task == 1 && P.st.pose_active && !P.st.random_base;
  (P.scene.env_box ? (g ? launch_reset_e1g1 : launch_reset_e1g0) : (g ? launch_reset_e0g1 : launch_reset_e0g0))(P, true, nullptr, nullptr, first_env, sm_count, s);
}
void launch_observe(const Params& P, cudaStream_t s) { (P.scene.env_box ? launch_observe_e1 : launch_observe_e0)(P, s); }

// The reset-pose table (Params::nprims / nextra).  q == nullptr: the generated constants of pg_neutral.h (panda.py:67).
// Otherwise the same quantities at the joint values q[7] a benchmark-scene builder installs (reach_ao.py:322, 343, 356,
// 424), by the fp32 forward kinematics of the step kernel -- 1e-7 m from the fp64 oracle, which only matters for a goal
// sample that lands within that distance of its rejection threshold.
struct NeutralQ { double q[7]; int custom; };
__global__ void k_neutral_table(const Params P, const NeutralQ nq) {
  const int k = threadIdx.x;
  if (!nq.custom) {
    if (k < pgm::kNumNeutralPrim) P.nprims[k] = pgm::kNeutralPrims[k];
    if (k < 3) { P.nextra[k] = pgm::kNeutralEE[k]; P.nextra[3 + k] = pgm::kNeutralLink0[k]; }
    if (k < 7) P.nextra[6 + k] = pgm::kNeutralQ[k];
    return;
  }
  if (k >= pgm::kNumNeutralPrim) return;
  float cq[7], sq[7];
  for (int d = 0; d < 7; d++) { cq[d] = (float)cos(nq.q[d]); sq[d] = (float)sin(nq.q[d]); }
  Frame F[8];
  F[0] = base_frame(v3(0.f, 0.f, 0.f));
  F[1] = frame_step<1>(F[0], cq[0], sq[0]); F[2] = frame_step<2>(F[1], cq[1], sq[1]); F[3] = frame_step<3>(F[2], cq[2], sq[2]);
  F[4] = frame_step<4>(F[3], cq[3], sq[3]); F[5] = frame_step<5>(F[4], cq[4], sq[4]); F[6] = frame_step<6>(F[5], cq[5], sq[5]);
  F[7] = frame_step<7>(F[6], cq[6], sq[6]);
  pgm::NeutralPrim np;
  if (k < pgm::kNumPrim) {
    const PrimRT& pr = kPrimRT[k];
    const V3 c = frame_point(F[pr.frame], pr.x, pr.y, pr.z), a = frame_dir(F[pr.frame], pr.ax, pr.ay, pr.az);
    np.type = pr.type; np.c[0] = c.x; np.c[1] = c.y; np.c[2] = c.z; np.a[0] = a.x; np.a[1] = a.y; np.a[2] = a.z; np.r = pr.r; np.h = pr.h;
  } else np = pgm::kNeutralPrims[k];  // panda_link0: fixed to the base
  P.nprims[k] = np;
  if (k == 0) {
    const V3 ee = frame_point(F[7], pgm::kEEx, pgm::kEEy, pgm::kEEz);
    const V3 l0 = frame_point(F[1], pgm::Link<1>::cx, pgm::Link<1>::cy, pgm::Link<1>::cz);
    P.nextra[0] = ee.x; P.nextra[1] = ee.y; P.nextra[2] = ee.z; P.nextra[3] = l0.x; P.nextra[4] = l0.y; P.nextra[5] = l0.z;
    for (int d = 0; d < 7; d++) P.nextra[6 + d] = nq.q[d];
  }
}
void launch_neutral_table(const Params& P, const double* q7, cudaStream_t s) {
  NeutralQ nq;
  nq.custom = q7 != nullptr;
  for (int d = 0; d < 7; d++) nq.q[d] = q7 ? q7[d] : 0.0;
  k_neutral_table<<<1, 64, 0, s>>>(P, nq);
}
}  // namespace pg
