#!/bin/bash
# compute-sanitizer over the small-size GPU tests (racecheck / memcheck / initcheck); summaries -> gpurun_out/
# usage (on the GPU box): bash tools/sanitize.sh
set -u
OUT=gpurun_out/sanitize
mkdir -p $OUT
CS=/usr/local/cuda/bin/compute-sanitizer
SEL='forward or backward_parity or sort or densify or adam'
FILES="tests/test_gpu_forward.py tests/test_gpu_backward.py tests/test_gpu_sort_adam_densify.py"
for tool in memcheck racecheck; do
  timeout 420 $CS --tool $tool --print-limit 20 --log-file $OUT/$tool.log \
    python -m pytest $FILES -x -q -m gpu -k "$SEL" -p no:cacheprovider > $OUT/$tool.pytest.txt 2>&1
  echo "$tool rc=$?" >> $OUT/summary.txt
  tail -3 $OUT/$tool.pytest.txt >> $OUT/summary.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard" $OUT/$tool.log | tail -5 >> $OUT/summary.txt
done
# initcheck on one C1 frame (fwd + bwd + Adam through smoke())
timeout 240 $CS --tool initcheck --print-limit 20 --log-file $OUT/initcheck.log python __graft_entry__.py --smoke > $OUT/initcheck.run.txt 2>&1
echo "initcheck rc=$?" >> $OUT/summary.txt
grep -E "ERROR SUMMARY" $OUT/initcheck.log | tail -2 >> $OUT/summary.txt
cat $OUT/summary.txt
