#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export INKSAM_WINDOW_ATTN=${WV:-5}; export INKSAM_WINDOW_KPOLY=${WK:-16}
timeout 300 python tools/attn_one.py window 3 > gpurun_out/attn_one.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_window -s 1 -c 1 -o gpurun_out/prof_winx python tools/attn_one.py window 3 > gpurun_out/ncu_winx.log 2>&1
echo "ncu win4 rc=$?"
