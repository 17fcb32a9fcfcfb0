#!/bin/bash
# Same-box A/B of two builds of the library: tools/_bin/old (a `git archive` of an earlier commit, built in place) against the
# working tree, alternating, main bench line only.  usage: tools/ab_bench.sh [rounds]
R=${1:-2}
for i in $(seq $R); do
  for side in old new; do
    if [ $side = old ]; then d=tools/_bin/old; else d=.; fi
    (cd $d && python bench.py --steps 100 --warmup 10 --no-extra --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$side', round(d['ms_per_step'],4), 'ms/step', 'e2e', round(d['e2e']['value']/1e6,2), 'M/s')")
  done
done
