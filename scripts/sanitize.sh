#!/bin/bash
# compute-sanitizer over a small-shape pass of every kernel family (smoke(): per-cell path, fused + staged
# network render on tcgen05, Frequency network, one training step).  ONE tool per gpurun call
# (B200_PROFILING.md): bash scripts/sanitize.sh memcheck|racecheck|synccheck  -> gpurun_out/sanitize_<tool>.txt
TOOL=${1:-memcheck}
O=gpurun_out
mkdir -p $O
python -c 'import __graft_entry__ as g; g.smoke()' > $O/sanitize_plain.txt 2>&1 || { echo "plain run failed"; tail -5 $O/sanitize_plain.txt; exit 1; }
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 --error-exitcode 3 \
  python -c 'import __graft_entry__ as g; g.smoke()' > $O/sanitize_$TOOL.txt 2>&1
echo "compute-sanitizer --tool $TOOL rc=$?" | tee -a $O/sanitize_$TOOL.txt
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|smoke [0-9] ok|Error|error" $O/sanitize_$TOOL.txt | head -20
