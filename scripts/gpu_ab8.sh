#!/bin/bash
# A/B of the brick-ordered occupancy bits in the march
O=gpurun_out/${1:-ab8}
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_grid_update.py tests/test_gpu_batch.py tests/test_zz_no_alloc.py tests/test_compat_cpp.py -x -q -m gpu > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee $O/rc.txt
RTXNERF_MARCH_FLAT=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q -m gpu > $O/pytest_flat.log 2>&1; echo "pytest(flat) rc=$?" | tee -a $O/rc.txt
for b in 0 1 0 1; do
  echo "== BRICKS=$b" >> $O/ab.txt
  RTXNERF_MARCH_BRICKS=$b timeout 200 python scripts/stage_bench.py 2>&1 | grep -iE "march" >> $O/ab.txt
  RTXNERF_MARCH_BRICKS=$b timeout 300 python scripts/stage_bench.py --G 512 --W 1920 --H 1080 --hmax 64 --occ 0.1 2>&1 | grep -iE "march" >> $O/ab.txt
done
cat $O/ab.txt; tail -n 3 $O/pytest.log $O/pytest_flat.log
