cp rtx_nerf_b200/librtxnerf_b200.so /tmp/lib_orig.so
for v in orig 4 64 orig 64; do
  if [ $v = orig ]; then cp /tmp/lib_orig.so rtx_nerf_b200/librtxnerf_b200.so; else cp build_variants/lib_$v.so rtx_nerf_b200/librtxnerf_b200.so; fi
  echo "== variant $v"; timeout 120 python scripts/train_bench.py --steps 8 | grep -E "net_bwd|step total"
  timeout 120 python scripts/batch_bench.py --reps 10 | grep "group=32)"
done
cp /tmp/lib_orig.so rtx_nerf_b200/librtxnerf_b200.so
