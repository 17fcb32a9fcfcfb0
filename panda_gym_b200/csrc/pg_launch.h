// pg_launch.h -- host-callable launchers of the kernels, one translation unit per kernel family (pg_tu_*.cu) so that the
// library builds in parallel and a change to one kernel recompiles one file.  pg_api.cu (the C ABI) only calls these.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pg {

struct Params;
struct HerParams;
struct HerBatch;
struct Camera {          // built on the host (pg_api.cu), Bullet's view / projection conventions
  float eye[3], fwd[3], right[3], up[3];
  float tan_half_fov, aspect, near_plane, far_plane;
  int width, height;
};

// pg_tu_step.cu
void launch_step(const Params& P, const float* actions, cudaStream_t s);
void launch_debug_substeps(int n, float* q, float* qd, const float* tgt, const float* f, int n_sub, int velocity_mode, cudaStream_t s);
cudaError_t step_kernel_attrs(int* out6);
// pg_tu_step_envbox.cu
void launch_step_envbox(const Params& P, const float* actions, cudaStream_t s);
// pg_tu_obj.cu
void launch_step_obj(const Params& P, const float* actions, cudaStream_t s);
// pg_tu_reset.cu
void launch_reset_all(const Params& P, const uint8_t* mask, const uint64_t* seeds, uint64_t first_env, cudaStream_t s);
void launch_reset_list(const Params& P, uint64_t first_env, int sm_count, cudaStream_t s);
void launch_observe(const Params& P, cudaStream_t s);
void launch_neutral_table(const Params& P, const double* q7 /* nullptr = panda.py:67 */, cudaStream_t s);
// pg_tu_reset_e{0,1}g{0,1}.cu: e = per-environment cuboid sizes, g = general start-pose randomizer; list = the auto-reset list
#define PG_DECL_RESET(tag) void launch_reset_##tag(const Params& P, bool list, const uint8_t* mask, const uint64_t* seeds, uint64_t first_env, int sm_count, cudaStream_t s);
PG_DECL_RESET(e0g0) PG_DECL_RESET(e0g1) PG_DECL_RESET(e1g0) PG_DECL_RESET(e1g1)
#undef PG_DECL_RESET
void launch_observe_e0(const Params& P, cudaStream_t s);
void launch_observe_e1(const Params& P, cudaStream_t s);
// pg_tu_misc.cu
void launch_mirror_obstacles(const Params& P, cudaStream_t s);
void launch_reward(const Params& P, const float* ag, const float* dg, float* out, int64_t n, int src, cudaStream_t s);
void launch_random_actions(float* out, int64_t m, uint64_t seed, uint64_t counter, cudaStream_t s);
void launch_eval_metrics(const Params& P, cudaStream_t s);
void launch_manipulability(int n, const float* q, float* out, cudaStream_t s);
void launch_debug_minv(int n, const float* q, float* out, cudaStream_t s);
void launch_debug_ik(int n, const float* q, const float* tgt, float bx, float by, float bz, float* out, cudaStream_t s);
void launch_fma_peak(int blocks, float* out, int iters, cudaStream_t s);

// pg_tu_render.cu
void launch_render(const Params& P, const Camera& cam, int env, uint8_t* out, cudaStream_t s);
// pg_tu_her.cu
void launch_her_add(const HerParams& H, const float* obs, const float* ag, const float* dg, const float* nobs, const float* nag,
                    const float* ndg, const float* act, const float* rew, const uint8_t* done, const uint8_t* timeout, cudaStream_t s);
void launch_her_compact(const HerParams& H, int total, cudaStream_t s);
void launch_her_indices(const HerParams& H, int B, int nbv, bool exact, uint64_t seed, uint64_t counter, cudaStream_t s);
void launch_her_gather(const HerParams& H, const Params& P, const HerBatch& O, int B, int nbv, int reward_env, int sm_count, cudaStream_t s);

}  // namespace pg
