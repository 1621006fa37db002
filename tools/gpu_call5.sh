#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/attn_one.py global 3 > gpurun_out/attn_one.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_global2 -s 1 -c 1 -o gpurun_out/prof_glob python tools/attn_one.py global 3 > gpurun_out/ncu_glob.log 2>&1
echo "ncu glob rc=$?"
