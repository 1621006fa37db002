#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/attn_one.py window 3 > gpurun_out/attn_one.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_window3 -s 1 -c 1 -o gpurun_out/prof_win3 python tools/attn_one.py window 3 > gpurun_out/ncu_win3.log 2>&1
echo "ncu win3 rc=$?"
INKSAM_WINDOW_ATTN=2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_window2 -s 1 -c 1 -o gpurun_out/prof_win2 python tools/attn_one.py window 3 > gpurun_out/ncu_win2.log 2>&1
echo "ncu win2 rc=$?"
ls -la gpurun_out/*.ncu-rep
