#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_attention.py -m gpu -q -s --tb=short -p no:cacheprovider -x -k "global or window_attention" > gpurun_out/attn_tests.log 2>&1; echo "attn tests rc=$?"; tail -25 gpurun_out/attn_tests.log
timeout 600 python tools/attn_bench.py > gpurun_out/attn_bench.log 2>&1; echo "attn bench rc=$?"; cat gpurun_out/attn_bench.log
