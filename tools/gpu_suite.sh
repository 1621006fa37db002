#!/bin/bash
# Runs -m gpu tests in separate processes under a timeout (a deadlocked kernel cannot hang the whole call).
# usage: tools/gpu_suite.sh [file[:k-expression] ...]   logs -> gpurun_out/suite_<n>_<file>.log
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
items=("$@")
[ ${#items[@]} -eq 0 ] && items=(test_gpu_gemm test_gpu_ops test_gpu_attention:global test_gpu_attention:window test_gpu_e2e)
rc_all=0; i=0
for it in "${items[@]}"; do
  f=${it%%:*}; k=""; [[ "$it" == *:* ]] && k=${it#*:}
  log=gpurun_out/suite_${i}_$f.log; i=$((i+1))
  if [ -n "$k" ]; then
    timeout 900 python -m pytest tests/$f.py -m gpu -k "$k" -q -s --tb=short -p no:cacheprovider > $log 2>&1
  else
    timeout 900 python -m pytest tests/$f.py -m gpu -q -s --tb=short -p no:cacheprovider > $log 2>&1
  fi
  rc=$?
  echo "== $it rc=$rc"; grep -E "passed|failed|error|Error|rel-RMS|FAILED|assert " $log | tail -n 30
  [ $rc -ne 0 ] && rc_all=1
done
exit $rc_all
