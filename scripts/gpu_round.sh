#!/bin/bash
# One gpurun call: gpu tests, smoke, contract bench (+ reference arm), stage benches, ncu launch list, ncu --set full
# captures, SASS mnemonic counts, roofline-traffic file.   Usage: gpurun --timeout 2400 -- 'bash scripts/gpu_round.sh TAG [noprof]'
TAG=${1:-x}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/rc.txt
timeout 300 python scripts/train_bench.py > $O/train_bench.txt 2>&1
timeout 300 python scripts/ref_kernels_bench.py --json $O/ref_kernels.json > $O/ref_kernels.txt 2>&1; echo "ref-kernels rc=$?" | tee -a $O/rc.txt
timeout 300 python scripts/freqnet_bench.py > $O/freqnet_bench.txt 2>&1
timeout 300 python scripts/batch_bench.py > $O/batch_bench.txt 2>&1; echo "batch-bench rc=$?" | tee -a $O/rc.txt
if [ "$2" != "noprof" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu --no-config5 --spinup 0 > $O/ncu_launch.log 2>&1; echo "ncu-list rc=$?" | tee -a $O/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -s 8 -c 4 -o $O/view_full -f \
  python bench.py --steps 2 --warmup 3 --no-cpu --no-train --no-refnet --no-config5 --spinup 0 > $O/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $O/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -s 12 -c 4 -o $O/view_staged_full -f \
  python bench.py --steps 2 --warmup 3 --no-cpu --no-train --no-refnet --no-config5 --spinup 0 --staged > $O/ncu_staged.log 2>&1; echo "ncu-staged rc=$?" | tee -a $O/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -s 24 -c 14 -o $O/train_full -f \
  python scripts/train_bench.py --steps 1 > $O/ncu_train.log 2>&1; echo "ncu-train rc=$?" | tee -a $O/rc.txt
for r in $O/*.ncu-rep; do python scripts/ncu_summary.py $r > ${r%.ncu-rep}.txt 2>/dev/null; done
python scripts/update_traffic.py --round $TAG "march=$O/view_full.ncu-rep:march_onepass" "sample+network+composite=$O/view_full.ncu-rep:rays_roles_kernel<(\(bool\))?1," \
  "sampler=$O/view_staged_full.ncu-rep:sampler_kernel" "network=$O/view_staged_full.ncu-rep:rays_roles_kernel<(\(bool\))?0," \
  "composite=$O/view_staged_full.ncu-rep:volrender_kernel" > $O/traffic.log 2>&1; echo "traffic rc=$?" | tee -a $O/rc.txt
cp profiles/ncu_traffic.json $O/ncu_traffic.json
# gpurun merges at most 64 MiB back: drop the largest reports until the directory fits
while [ $(du -sm gpurun_out | cut -f1) -ge 60 ]; do big=$(ls -S $O/*.ncu-rep 2>/dev/null | head -1); [ -z "$big" ] && break; rm -f $big; done
fi
tail -3 $O/pytest.log; cat $O/smoke.log | tail -4; cat $O/bench.json | cut -c1-400; cat $O/train_bench.txt; tail -3 $O/freqnet_bench.txt; cat $O/rc.txt; cat $O/traffic.log | cut -c1-600
