#!/bin/bash
# ncu --set full of the CTA-pair GEMM (perf group of tools/gemm_check.py) to find out why it is slow.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/gemm_check.py perf_bf16_cg2 > gpurun_out/cg2_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 3 -c 2 -o gpurun_out/prof_cg2 python tools/gemm_check.py perf_bf16_cg2 > gpurun_out/ncu_cg2.log 2>&1
echo "ncu cg2 rc=$?"
python tools/gemm_check.py perf_bf16_cg1 > gpurun_out/cg1_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 3 -c 2 -o gpurun_out/prof_cg1 python tools/gemm_check.py perf_bf16_cg1 > gpurun_out/ncu_cg1.log 2>&1
echo "ncu cg1 rc=$?"
