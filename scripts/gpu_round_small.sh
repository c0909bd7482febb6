O=gpurun_out/r01_i; mkdir -p $O
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --spinup 0 > $O/ncu_launch.log 2>&1; echo "ncu-list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:batch_ -s 30 -c 3 -o $O/batch_full -f python scripts/batch_bench.py --reps 3 > $O/ncu_batch.log 2>&1; echo "ncu-batch rc=$?"
python -c "import json;d=json.load(open('$O/bench.json'));print(json.dumps(d['train_step'],indent=1)); print(d['value'], d['e2e'], d['gpu_launches'])"
