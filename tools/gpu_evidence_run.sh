#!/bin/bash
# Round-2 evidence run: the whole -m gpu suite, the bench line (+ reference arm), configs 3 and 5, the DRAM traffic of one
# decode pass, the ncu launch list of one bench step and --set full captures of the top kernels.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tools/gpu_suite.sh test_gpu_gemm test_gpu_gemm_stress test_gpu_ops test_gpu_attention test_gpu_ln_fold test_gpu_refine \
    test_gpu_e2e test_gpu_config1 test_properties test_zz_no_alloc
echo "suite rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
timeout 600 python bench.py --config 3 --steps 10 > gpurun_out/config3.json 2> gpurun_out/config3.err; echo "config3 rc=$?"; cat gpurun_out/config3.json
timeout 900 python bench.py --config 5 > gpurun_out/config5.json 2> gpurun_out/config5.err; echo "config5 rc=$?"; cat gpurun_out/config5.json
timeout 300 python tools/decode_pass.py > gpurun_out/decode_pass.log 2>&1 && cat gpurun_out/decode_pass.log &&
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --profile-from-start off \
    --csv --log-file gpurun_out/dec_traffic.csv python tools/decode_pass.py > gpurun_out/ncu_dec_traffic.log 2>&1
echo "ncu decoder traffic rc=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 460 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 60 -c 8 -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "ncu gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_ -s 6 -c 3 -o gpurun_out/prof_attn $CMD > gpurun_out/ncu_attn.log 2>&1
echo "ncu attn rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"postprocess|decoder_" --profile-from-start off -o gpurun_out/prof_dec python tools/decode_pass.py > gpurun_out/ncu_dec.log 2>&1
echo "ncu decoder rc=$?"
ls -la gpurun_out | head -50
