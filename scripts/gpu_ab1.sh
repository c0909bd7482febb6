#!/bin/bash
# round-2 step 1: decoupled-role kernels against the round-1 kernels + the network / full-size / training tests
O=gpurun_out/r02_a
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 300 python scripts/ab_fused.py --tag new --out $O/px_new.npy > $O/ab_new.txt 2>&1; echo "ab new rc=$?" | tee -a $O/rc.txt
RTXNERF_FUSED_V1=1 RTXNERF_NETR_VERSION=2 timeout 300 python scripts/ab_fused.py --tag v1 --out $O/px_v1.npy > $O/ab_v1.txt 2>&1; echo "ab v1 rc=$?" | tee -a $O/rc.txt
python - <<'PY' > $O/ab_diff.txt 2>&1
import numpy as np
a=np.load("gpurun_out/r02_a/px_new.npy"); b=np.load("gpurun_out/r02_a/px_v1.npy")
print("new vs v1 pixels: max abs diff %.3g, identical %s" % (np.abs(a-b).max(), np.array_equal(a,b)))
PY
timeout 900 python -m pytest tests/test_gpu_network.py tests/test_gpu_full_size.py tests/test_gpu_training.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
cat $O/ab_new.txt $O/ab_v1.txt $O/ab_diff.txt; tail -15 $O/pytest.log; cat $O/rc.txt
