#!/bin/bash
O=gpurun_out/r02_v
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_freqnet.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
grep -E "passed|failed|^E" $O/pytest.log | head -20
