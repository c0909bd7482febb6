#!/bin/bash
O=gpurun_out/r02_n
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
timeout 300 python scripts/train_bench.py > $O/train_bench.txt 2>&1
tail -5 $O/pytest.log; tail -2 $O/smoke.log; cut -c1-1500 $O/bench.json; cat $O/train_bench.txt; cat $O/rc.txt
