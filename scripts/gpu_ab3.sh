#!/bin/bash
O=gpurun_out/r02_u
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_training.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -3 $O/pytest.log
for i in 1 2; do
RTXNERF_MLPB_V1=1 timeout 300 python scripts/train_bench.py 2>&1 | grep -E "net_bwd|step total" | sed 's/^/v1  /' | tee -a $O/ab.txt
timeout 300 python scripts/train_bench.py 2>&1 | grep -E "net_bwd|step total" | sed 's/^/new /' | tee -a $O/ab.txt
done
