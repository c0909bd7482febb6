#!/bin/bash
# One gpurun --gpus 8 call: the contract bench under torchrun at N = 8, 4, 2 and the reference arm at N = 8.
TAG=${1:-x}
O=gpurun_out/$TAG
mkdir -p $O
for N in 8 4 2; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2950$N \
    bench.py --gpus $N --steps 20 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "N=$N rc=$?" | tee -a $O/rc.txt
  tail -1 $O/bench_n$N.json | cut -c1-250
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > $O/bench_ref_n8.json 2> $O/bench_ref_n8.err; echo "ref N=8 rc=$?" | tee -a $O/rc.txt
tail -1 $O/bench_ref_n8.json | cut -c1-200
