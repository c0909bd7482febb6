cp rtx_nerf_b200/librtxnerf_b200.so /tmp/lib_orig.so
for v in orig 8 40 orig 40; do
  if [ $v = orig ]; then cp /tmp/lib_orig.so rtx_nerf_b200/librtxnerf_b200.so; else cp build_variants/lib_$v.so rtx_nerf_b200/librtxnerf_b200.so; fi
  echo "== variant $v"; timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu --no-train --no-refnet | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value']/1e6, d['ms_per_step'], d['stages']['sample+network+composite']['ms'])"
done
cp /tmp/lib_orig.so rtx_nerf_b200/librtxnerf_b200.so
timeout 200 python scripts/train_bench.py --steps 8 | grep -E "net_bwd|step total"
timeout 300 python -m pytest tests/test_gpu_training.py -x -q | tail -1
