// pg_step_body.cuh -- body of the ReachAO / Reach step kernel (RobotTaskEnv.step, core.py:352-368), included once per
// translation unit that instantiates it: pg_tu_step.cu (PG_ENVBOX 0: obstacle shapes straight from the kernel parameters, the
// default) and pg_tu_step_envbox.cu (PG_ENVBOX 1: per-environment cuboid sizes, scenario reachao_rand_shape).
// The includer defines PG_STEP_KERNEL_NAME and opens namespace pg.
constexpr int kStepSmem = 57 * kStepBlock * (int)sizeof(float);  // tgt 7 | q 7 | q0 7 | slack 9 | ctr0 27

// One environment per thread, one 448-thread block per SM.
__global__ void PG_STEP_BOUNDS PG_STEP_KERNEL_NAME(const Params P, const float* __restrict__ actions) {
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

