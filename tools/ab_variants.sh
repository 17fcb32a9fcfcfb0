#!/bin/bash
# Same-box A/B of library variants (tools/build_variants.py) against the default build: main bench line, alternating.
# usage: tools/ab_variants.sh [rounds] [bench args...]
R=${1:-2}; shift
for i in $(seq $R); do
  for lib in panda_gym_b200/variants/*.so panda_gym_b200/libpanda_b200.so; do
    PANDA_B200_LIB=$PWD/$lib python bench.py --steps 200 --warmup 10 --no-extra --no-cpu-baseline "$@" 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$(basename $lib)', round(d['ms_per_step'],4), 'ms/step median', round(d['step_ms']['median'],4), 'min', round(d['step_ms']['min'],4))"
  done
done
