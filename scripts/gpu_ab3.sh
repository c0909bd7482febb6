#!/bin/bash
O=gpurun_out/r02_o
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_training.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -5 $O/pytest.log
timeout 300 python scripts/train_bench.py 2>&1 | tee $O/train_new.txt | grep -E "net_bwd|step total"
RTXNERF_SCATTER_V1=1 timeout 300 python scripts/train_bench.py 2>&1 | tee $O/train_v1.txt | grep -E "net_bwd|step total"
