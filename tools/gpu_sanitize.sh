#!/bin/bash
# compute-sanitizer pass over the op-level GPU tests (memcheck + racecheck on every kernel).
# Usage (from the repo root, under gpurun):  bash tools/gpu_sanitize.sh [tag] [pytest -k expression]
# Summaries land in gpurun_out/<tag>_sanitizer_{memcheck,racecheck}.txt (copy into profiles/).
TAG=${1:-r02}
SEL=${2:-"not train and not fullsize"}
O=gpurun_out
mkdir -p $O
for TOOL in memcheck racecheck; do
  LOG=$O/${TAG}_sanitizer_${TOOL}.txt
  echo "== compute-sanitizer --tool $TOOL: pytest tests/test_gpu_ops.py tests/test_gpu_linear.py -k '$SEL'" > $LOG
  timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 --error-exitcode 86 \
      python -m pytest tests/test_gpu_ops.py tests/test_gpu_linear.py -x -q -m gpu -k "$SEL" -p no:cacheprovider >> $LOG.full 2>&1
  RC=$?
  echo "exit code: $RC (86 = sanitizer errors, 124 = timeout)" >> $LOG
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" $LOG.full | tail -20 >> $LOG
  grep -E "=========( |	)(Invalid|Race|Error|Uninitialized|Program hit)" $LOG.full | sort | uniq -c | sort -rn | head -40 >> $LOG
  tail -c 20000 $LOG.full > $LOG.tail; rm -f $LOG.full
  cat $LOG
done
