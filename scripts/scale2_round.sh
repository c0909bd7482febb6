#!/bin/bash
# One gpurun --gpus 2 call: the contract bench under torchrun at N = 2 (both arms).
TAG=${1:-x}
O=gpurun_out/$TAG
mkdir -p $O
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 \
  bench.py --gpus 2 --steps 20 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; echo "N=2 rc=$?" | tee -a $O/rc.txt
tail -1 $O/bench_n2.json | cut -c1-250
python -c "import json;d=json.loads(open('$O/bench_n2.json').read().strip().splitlines()[-1]);print(json.dumps(d['train_step'],indent=1))"
tail -5 $O/bench_n2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/bench_ref_n2.json 2> $O/bench_ref_n2.err; echo "ref N=2 rc=$?" | tee -a $O/rc.txt
tail -1 $O/bench_ref_n2.json | cut -c1-200
