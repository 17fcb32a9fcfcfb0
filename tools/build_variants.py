"""Build A/B variants of the library (extra -D flags) into panda_gym_b200/variants/ for tools/time_variants.py.
usage: build_variants.py name1="-DX=1 -DY=2" name2="..." """
import os, subprocess, sys
from concurrent.futures import ThreadPoolExecutor
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "panda_gym_b200", "csrc")
OUT = os.path.join(ROOT, "panda_gym_b200", "variants")
os.makedirs(OUT, exist_ok=True)
BASE = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC"]
units = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
def build(arg):
    name, flags = arg.split("=", 1)
    d = os.path.join(OUT, "obj_" + name); os.makedirs(d, exist_ok=True)
    def cu(u):
        o = os.path.join(d, u[:-3] + ".o")
        subprocess.run(BASE + flags.split() + ["-c", "-o", o, os.path.join(CSRC, u)], check=True, cwd=CSRC)
        return o
    with ThreadPoolExecutor(4) as ex:
        objs = list(ex.map(cu, units))
    subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", os.path.join(OUT, "libpanda_b200_%s.so" % name)] + objs, check=True)
    print("built", name)
with ThreadPoolExecutor(2) as ex:
    list(ex.map(build, sys.argv[1:]))
