#!/bin/bash
# A/B of the scatter's run cap
O=gpurun_out/${1:-ab6}
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_training.py -x -q -m gpu > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee $O/rc.txt
RTXNERF_HB_MAX_RUN=4 timeout 600 python -m pytest tests/test_gpu_training.py -x -q -m gpu > $O/pytest4.log 2>&1; echo "pytest(4) rc=$?" | tee -a $O/rc.txt
for mr in 32 16 8 4 2 1; do
  echo "== HB_MAX_RUN=$mr" >> $O/ab.txt
  RTXNERF_HB_MAX_RUN=$mr timeout 200 python scripts/train_bench.py 2>&1 | grep -E "net_bwd|step total" >> $O/ab.txt
done
cat $O/ab.txt; tail -3 $O/pytest.log $O/pytest4.log
