#!/bin/bash
O=gpurun_out/r02_x
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_freqnet.py -x -q -s > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
grep -E "passed|failed|^E  |PSNR" $O/pytest.log | head
timeout 900 python bench.py --no-cpu --no-config5 --steps 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_x/bench.json"))
print(json.dumps(d["reference_network"]["train_step"], indent=1))
PY
