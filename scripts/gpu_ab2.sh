#!/bin/bash
O=gpurun_out/r02_f
mkdir -p $O
for cfg in "3 0" "3 1" "3 2" "4 1"; do
  set -- $cfg
  RTXNERF_RR_GROUPS=$1 RTXNERF_RR_DEV_SKIP=$2 timeout 300 python scripts/ab_fused.py --tag "eg$1-skip$2" 2>&1 | grep -E "eg|fused |vs staged" | tee -a $O/sweep.txt
done
