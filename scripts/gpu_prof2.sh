#!/bin/bash
# ncu --set full captures (with source) of the training backward kernels and the march
O=gpurun_out/${1:-prof2}
mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"hashgrid_backward|mlp_backward" -s 4 -c 2 -o $O/train_bwd -f \
  python scripts/train_bench.py --steps 1 > $O/ncu_train.log 2>&1; echo "ncu-train rc=$?" | tee $O/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:march_onepass -s 2 -c 1 -o $O/march -f \
  python scripts/ab_fused.py > $O/ncu_march.log 2>&1; echo "ncu-march rc=$?" | tee -a $O/rc.txt
for r in $O/*.ncu-rep; do python scripts/ncu_summary.py $r > ${r%.ncu-rep}.txt 2>/dev/null; done
