#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tools/gpu_suite.sh test_gpu_gemm test_gpu_gemm_stress test_gpu_ln_fold
timeout 300 python tools/fold_perf.py 2>&1 | tail -8
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pre.json 2> gpurun_out/bench_pre.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_pre.json"))
print("value", round(d["value"],2), "e2e", round(d["e2e"]["value"],2), "attn share", round(d["encoder"]["attention_share_of_step"],4), "enc ms", round(d["encoder"]["ms_per_batch"],3), "gemm frac", round(d["roofline"]["frac"],4), "avg gemm ms", round(d["roofline"]["avg_launch_ms"],5), "parity", round(d["parity"]["min_iou"],4), "clk", d["clocks"]["sm_mhz"])
PY
