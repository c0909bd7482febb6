#!/bin/bash
O=gpurun_out/r02_r
mkdir -p $O
for g in 2 4 8 16; do
  RTXNERF_RR_GROUPS_PER_CTA=$g timeout 300 python scripts/ab_fused.py --tag "gpc$g" 2>&1 | grep -E "gpc|fused |network" | tee -a $O/sweep.txt
done
