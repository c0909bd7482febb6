#!/bin/bash
# A/B of the four-tiles-in-flight mlp_backward kernel
O=gpurun_out/${1:-ab7}
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_training.py tests/test_gpu_analytic.py -x -q -m gpu > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee $O/rc.txt
for v in 0 1 0 1; do
  echo "== MLPB_SLOTS=$v" >> $O/ab.txt
  RTXNERF_MLPB_SLOTS=$v timeout 200 python scripts/train_bench.py 2>&1 | grep -E "net_bwd|step total" >> $O/ab.txt
done
timeout 200 python scripts/mlp_bench.py > $O/mlp_bench.txt 2>&1
cat $O/ab.txt; tail -n 5 $O/pytest.log; tail -n 8 $O/mlp_bench.txt
