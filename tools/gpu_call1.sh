#!/bin/bash
# Round-2 GPU call 1: the whole -m gpu suite (one process per file, under timeouts), the default bench line, the
# reference arm, and the DRAM traffic of one decode pass.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
tools/gpu_suite.sh test_gpu_gemm test_gpu_gemm_stress test_gpu_ops test_gpu_attention test_gpu_ln_fold test_gpu_refine \
    test_gpu_e2e test_gpu_config1 test_properties
echo "suite rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
timeout 300 python tools/decode_pass.py > gpurun_out/decode_pass.log 2>&1 && cat gpurun_out/decode_pass.log &&
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --profile-from-start off \
    --csv --log-file gpurun_out/dec_traffic.csv python tools/decode_pass.py > gpurun_out/ncu_dec_traffic.log 2>&1
echo "ncu decoder traffic rc=$?"
timeout 600 python bench.py --config 3 --steps 10 > gpurun_out/config3.json 2> gpurun_out/config3.err; echo "config3 rc=$?"; cat gpurun_out/config3.json
ls -la gpurun_out | head -40
