#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for v in 2 4 2 4; do
  INKSAM_WINDOW_ATTN=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_w$v.json 2> gpurun_out/bench_w$v.err
  echo "window v$v rc=$?"; python - <<PY
import json
d=json.load(open("gpurun_out/bench_w$v.json"))
print("v$v value", round(d["value"],2), "e2e", round(d["e2e"]["value"],2), "attn share", round(d["encoder"]["attention_share_of_step"],4), "enc ms", round(d["encoder"]["ms_per_batch"],3), "gemm frac", round(d["roofline"]["frac"],4), "parity", round(d["parity"]["min_iou"],4), "clk", d["clocks"]["sm_mhz"])
PY
done
