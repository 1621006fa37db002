#!/bin/bash
# Per-kernel duration, DRAM bytes and pipe utilisation of one decode pass (tools/decode_pass.py) -> gpurun_out/dec_kernels.csv
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 500 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active \
  --clock-control none --profile-from-start off --csv --log-file gpurun_out/dec_kernels.csv python tools/decode_pass.py "$@" > /dev/null 2>&1
echo "rc=$?"
python tools/decode_kernels.py gpurun_out/dec_kernels.csv
