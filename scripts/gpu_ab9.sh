#!/bin/bash
# which of the round's changes moved the bench line's training leg
O=gpurun_out/${1:-ab9}
mkdir -p $O
run() { echo "== $1" >> $O/ab.txt; env $1 timeout 300 python bench.py --no-cpu --no-refnet --no-config5 --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
t=d['train_step']; print('view ms', round(d['ms_per_step'],3), 'train per-view', round(t['ms_per_step'],3), 'store-fed', round(t['shuffled']['ms_per_step'],3))" >> $O/ab.txt; }
run "RTXNERF_HB_MAX_RUN=8"
run "RTXNERF_HB_MAX_RUN=32"
run "RTXNERF_HB_MAX_RUN=16"
run "RTXNERF_HB_MAX_RUN=32 RTXNERF_RR_NAMED_BARS=0"
run "RTXNERF_HB_MAX_RUN=32 RTXNERF_MARCH_BRICKS=0"
run "RTXNERF_HB_MAX_RUN=8"
cat $O/ab.txt
