#!/bin/bash
O=gpurun_out/r02_l
mkdir -p $O
python scripts/ab_fused.py --tag plain > $O/plain.txt 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:rays_roles -s 6 -c 1 -o $O/rr_full -f \
  python scripts/ab_fused.py --tag ncu > $O/ncu.log 2>&1; echo "ncu rc=$?" | tee -a $O/rc.txt
python scripts/ncu_summary.py $O/rr_full.ncu-rep > $O/rr_full.txt 2>/dev/null
cat $O/plain.txt; cat $O/rr_full.txt | head -40
