#!/bin/bash
# A/B of the named-barrier hand-offs in the rays kernel + one ncu --set full capture of the fused kernel
O=gpurun_out/${1:-ab5}
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_network.py tests/test_gpu_full_size.py tests/test_gpu_training.py -x -q -m gpu > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee $O/rc.txt
for nb in 0 1 0 1; do
  echo "== NAMED_BARS=$nb" >> $O/ab.txt
  RTXNERF_RR_NAMED_BARS=$nb timeout 200 python scripts/ab_fused.py >> $O/ab.txt 2>&1
done
for lay in 31 42 41; do
  echo "== NAMED_BARS=1 LAYOUT=$lay" >> $O/ab.txt
  RTXNERF_RR_LAYOUT=$lay timeout 200 python scripts/ab_fused.py >> $O/ab.txt 2>&1
done
timeout 200 python scripts/train_bench.py > $O/train_bench.txt 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:rays_roles -s 3 -c 1 -o $O/rr_full -f \
  python scripts/ab_fused.py > $O/ncu.log 2>&1; echo "ncu rc=$?" | tee -a $O/rc.txt
python scripts/ncu_summary.py $O/rr_full.ncu-rep > $O/rr_full.txt 2>/dev/null
tail -3 $O/pytest.log; grep -E "==|fused  |network\(st" $O/ab.txt; cat $O/train_bench.txt | tail -4
