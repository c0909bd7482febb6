#!/bin/bash
# One gpurun --gpus 8 call: two-device-thread test, the overlap check at N = 2, the contract bench at N = 8 and N = 2.
TAG=${1:-x}
O=gpurun_out/$TAG
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_multidevice.py -x -q -m gpu > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/check_overlap.py > $O/overlap.log 2>&1; echo "overlap rc=$?" | tee -a $O/rc.txt
grep "^rank" $O/overlap.log | head -4
for N in 8 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2950$N \
    bench.py --gpus $N --steps 20 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "N=$N rc=$?" | tee -a $O/rc.txt
  tail -1 $O/bench_n$N.json | cut -c1-200
done
