#!/bin/bash
O=gpurun_out/r02_q
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_analytic.py -x -q -s > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
grep -E "PSNR|passed|failed|^E" $O/pytest.log | head -20
