#!/bin/bash
# One GPU call: plain bench (N=1) + reference arm, then the ncu launch list of the same command and full captures of
# the top kernels (each ncu pass only after the same command has exited 0 without ncu).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 460 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 60 -c 8 -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "ncu gemm rc=$?"
$CMD > gpurun_out/plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_ -s 6 -c 3 -o gpurun_out/prof_attn $CMD > gpurun_out/ncu_attn.log 2>&1
echo "ncu attn rc=$?"
DCMD="python tools/decode_bench.py"
$DCMD > gpurun_out/decode_bench.log 2>&1 && cat gpurun_out/decode_bench.log &&
ncu --set full --clock-control none --import-source on -k regex:"postprocess|decoder_|layernorm" -s 40 -c 14 -o gpurun_out/prof_dec $DCMD > gpurun_out/ncu_dec.log 2>&1
echo "ncu decoder rc=$?"
ls -la gpurun_out
