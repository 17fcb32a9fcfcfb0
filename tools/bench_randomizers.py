"""ms per step of the start-pose randomisers and per-episode cuboid sizes at BASELINE size (65 536 environments, auto-reset,
random actions, 150-step pre-roll): what the fp64 inverse kinematics in the reset costs next to the plain stage."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from panda_gym_b200 import sim
from panda_gym_b200.config import TrainConfig
N, P, K = 65536, 120, 60
out = {"envs": N, "preroll": P, "steps": K, "cases": {}}
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for name, scenario, randomize, limit in [
        ("reachao3", "reachao3", False, 0), ("reachao3 + randomize_robot_pose (ik_goal_space)", "reachao3", True, 0),
        ("reachao3 + randomize_robot_pose, pose_retry_limit 100", "reachao3", True, 100),
        ("reachao1 + randomize_robot_pose (torus, z window)", "reachao1", True, 0),
        ("wang-3 + randomize_robot_pose (torus)", "wang-3", True, 0), ("wang-3 + randomize_robot_pose, pose_retry_limit 100", "wang-3", True, 100),
        ("wangexp-3 + randomize_robot_pose (random base)", "wangexp-3", True, 0),
        ("reachao_rand", "reachao_rand", False, 0), ("reachao_rand_shape", "reachao_rand_shape", False, 0),
        ("reachao_rand_start (ik_sphere), pose_retry_limit 100", "reachao_rand_start", False, 100)]:
    tc = TrainConfig(); tc.randomize_robot_pose = randomize
    b = sim.PandaBatch(sim.make_config(1, N, scenario=scenario, config=tc, max_episode_steps=100, auto_reset=True, pose_retry_limit=limit))
    b.seed(5)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); b.reset(); e1.record(); torch.cuda.synchronize()
    first_reset_ms = e0.elapsed_time(e1)
    for k in range(P):
        b.step(b.random_actions(5, k))
    acts = [b.random_actions(6, k) for k in range(K)]
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for k in range(K):
        flush.fill_(k & 255)
        ev[k][0].record(); b.step(acts[k]); ev[k][1].record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(c) for a, c in ev)
    st = b.view("stats").cpu().numpy()
    out["cases"][name] = {"full_reset_ms": first_reset_ms, "ms_per_step_mean": sum(ms) / K, "median": ms[K // 2], "max": ms[-1],
                          "env_steps_per_s": N * K / (sum(ms) * 1e-3), "episodes": float(st[0]), "reset_failures": float(st[6])}
    print("%-52s full reset %8.2f ms | step mean %.3f median %.3f max %.3f ms | reset failures %d / %d episodes" % (
        name, first_reset_ms, sum(ms) / K, ms[K // 2], ms[-1], st[6], st[0]), file=sys.stderr)
    b.close()
print(json.dumps(out))
