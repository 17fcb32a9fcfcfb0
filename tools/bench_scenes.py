"""ms per step of every benchmark scene (static box sets, pg_set_scene), 16 384 environments, auto-reset, random actions."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from panda_gym_b200 import sim, scenes
N, W, K = 16384, 5, 30
out = {"envs": N, "steps": K, "warmup": W, "scenes": {}}
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for name in ["reachao3"] + list(scenes.SCENES):
    cfg = sim.make_config(1, N, scenario=name, max_episode_steps=100, auto_reset=True)
    b = sim.PandaBatch(cfg); b.seed(11); b.reset()
    acts = [b.random_actions(11, k) for k in range(W + K)]
    for k in range(W):
        b.step(acts[k])
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for k in range(K):
        flush.fill_(k & 255)
        ev[k][0].record(); b.step(acts[W + k]); ev[k][1].record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(c) for a, c in ev) / K
    st = b.view("stats").cpu().numpy()
    out["scenes"][name] = {"boxes": 0 if name == "reachao3" else int(len(scenes.builtin_scene_boxes(name))), "ms_per_step": ms,
                           "env_steps_per_s": N / (ms * 1e-3), "episodes": float(st[0]), "collisions": float(st[2])}
    print("%-16s %2d boxes  %.3f ms/step  %.2e env-steps/s" % (name, out["scenes"][name]["boxes"], ms, N / (ms * 1e-3)), file=sys.stderr)
    b.close()
print(json.dumps(out))
