#!/bin/bash
# A/B of development knobs: bash scripts/ab_bench.sh "ENV1=.. ENV2=.." "ENVB=.." ...
for e in "$@"; do
  echo "== $e"
  env $e python bench.py --no-cpu --no-train --no-refnet --steps 10 $BENCH_ARGS 2>&1 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value %.1f Mrays/s  ms/step %.3f  e2e %.3f ms' % (d['value']/1e6, d['ms_per_step'], d['e2e']['ms_per_step']), {k:round(v['ms'],3) for k,v in d['stages'].items()})
"
done
