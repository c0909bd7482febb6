#!/bin/bash
O=gpurun_out/r02_n2
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_multidevice.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -3 $O/pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/check_overlap.py > $O/overlap.log 2>&1; echo "overlap rc=$?" | tee -a $O/rc.txt
grep "rank" $O/overlap.log | head
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench rc=$?" | tee -a $O/rc.txt
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_n2/bench_n2.json"))
print(d["value"], d["ms_per_step"], d["train_step"]["ms_per_step"], d["train_step"]["allreduce_ms"], d["train_step"]["allreduce_alone_ms"], d["train_step"]["shuffled"]["ms_per_step"])
PY
