"""Generate panda_gym_b200/data/benchmark_scenes.json (numbers only: centre, half extents, margin of every collision box
of every benchmark scene) from a panda_gym checkout's assets/scenarios with panda_gym_b200.scenes.load_scene_boxes.
usage: python tools/gen_scene_data.py /root/reference/panda_gym/assets/scenarios"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib.util
spec = importlib.util.spec_from_file_location("scenes", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "panda_gym_b200", "scenes.py"))
scenes = importlib.util.module_from_spec(spec); spec.loader.exec_module(scenes)
assets = sys.argv[1]
out = {}
for name, s in scenes.SCENES.items():
    if s["asset"] is None or s["asset"] in out:
        continue
    b = scenes.load_scene_boxes(assets, name)
    out[s["asset"]] = [[round(float(v), 9) for v in row] for row in b]
    print("%-16s %3d boxes  x [%.2f, %.2f]  y [%.2f, %.2f]  z [%.2f, %.2f]" % (s["asset"], len(b), (b[:, 0] - b[:, 3]).min(), (b[:, 0] + b[:, 3]).max(),
          (b[:, 1] - b[:, 4]).min(), (b[:, 1] + b[:, 4]).max(), (b[:, 2] - b[:, 5]).min(), (b[:, 2] + b[:, 5]).max()))
json.dump(out, open(os.path.join(os.path.dirname(scenes.__file__), "data", "benchmark_scenes.json"), "w"), indent=0)
