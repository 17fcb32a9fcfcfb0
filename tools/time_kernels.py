"""Time pg_step with the one-env-per-thread kernel (1) and the packed two-env-per-thread kernel (2), subprocess each.
Kernel 2 only exists in a -DPG_PAIR_KERNEL build:
  cd panda_gym_b200/csrc && mkdir -p ../variants && nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo \
     -DPG_PAIR_KERNEL -Xcompiler -fPIC -shared -o ../variants/libpanda_b200_pair.so pg_api.cu"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, os, torch
sys.path.insert(0, %r)
from panda_gym_b200 import sim
sc = sys.argv[1]
cfg = sim.make_config(1, 65536, scenario=sc, max_episode_steps=100, auto_reset=True)
b = sim.PandaBatch(cfg); b.seed(1234); b.reset()
acts = [b.random_actions(1234, k) for k in range(30)]
for k in range(5): b.step(acts[k])
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for k in range(5, 30): b.step(acts[k])
e1.record(); torch.cuda.synchronize()
print("%%.4f ms/step  %%.3e steps/s" %% (e0.elapsed_time(e1) / 25, 65536 * 25 / e0.elapsed_time(e1) * 1e3), b.view("stats").cpu().numpy()[:6])
''' % ROOT
scen = sys.argv[1:] or ["reachao3"]
for sc in scen:
    for k in ("1", "2"):
        env = dict(os.environ, PANDA_B200_STEP_KERNEL=k, PANDA_B200_LIB=os.path.join(ROOT, "panda_gym_b200", "variants", "libpanda_b200_pair.so"))
        r = subprocess.run([sys.executable, "-c", CHILD, sc], env=env, capture_output=True, text=True)
        print(sc, "kernel", k, r.stdout.strip(), r.stderr.strip()[-300:], flush=True)
