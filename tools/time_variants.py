"""Time k_step for each tuning variant in panda_gym_b200/variants (subprocess per library)."""
import glob, os, subprocess, sys
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
print("%%.4f ms/step  %%.3e steps/s" %% (e0.elapsed_time(e1) / 25, 65536 * 25 / e0.elapsed_time(e1) * 1e3), b.view("stats").cpu().numpy()[:4])
''' % ROOT
scen = sys.argv[1] if len(sys.argv) > 1 else "reachao3"
libs = sorted(glob.glob(os.path.join(ROOT, "panda_gym_b200", "variants", "*.so"))) + [os.path.join(ROOT, "panda_gym_b200", "libpanda_b200.so")]
for lib in libs:
    env = dict(os.environ, PANDA_B200_LIB=lib)
    r = subprocess.run([sys.executable, "-c", CHILD, scen], env=env, capture_output=True, text=True)
    print(os.path.basename(lib), r.stdout.strip(), r.stderr.strip()[-300:])
