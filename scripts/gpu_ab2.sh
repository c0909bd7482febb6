#!/bin/bash
O=gpurun_out/r02_k
mkdir -p $O
for cfg in 32 31 42; do
  RTXNERF_RR_LAYOUT=$cfg timeout 300 python scripts/ab_fused.py --tag "layout$cfg" 2>&1 | grep -E "layout|fused |network|vs staged" | tee -a $O/sweep.txt
done
timeout 900 python -m pytest tests/test_gpu_network.py tests/test_gpu_full_size.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -5 $O/pytest.log
