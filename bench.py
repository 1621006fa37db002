"""Benchmark of the SAM segmentation hot path on B200 (BASELINE.json: sketches/sec, 1024^2 ViT-H encode + all-box decode).

    python bench.py --gpus N --steps K --warmup W            # this implementation, one rank per GPU
    python bench.py --impl reference --steps K --warmup W    # the reference algorithm on the host cores (oracle port)

A step = one batch of 8 synthetic 1024x1024 sketches per GPU: ViT-H image encoder on the batch, then the mask
decoder + fused post-process on 16 boxes of every sketch (config 4 of BASELINE.json at per-step granularity;
config 2 is its encoder part).  Prints ONE JSON line (rank 0).
  value      sketches/s with inputs resident in HBM (CUDA events, max over ranks)
  e2e        sketches/s through the public API (SketchSegmenter.segment_batch) from pinned host images to host masks
  roofline   the dominant kernel (tcgen05 GEMM, every encoder instantiation): algorithmic 2*M*N*K per launch over the
             CUDA-event duration of those launches inside the timed steps, against the measured sustained bf16 peak
  cpu_baseline   the oracle (CPU restatement of the reference) timed on this box's host cores on a bounded sample
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH = 8           # sketches per step per GPU
BOXES = 16          # boxes per sketch
MODEL = "vit_h"
ENC_GFLOP = {"vit_h": 5641.8, "vit_l": 2837.0, "vit_b": 937.6}  # algorithmic, SURVEY.md 8(d)
METRIC = "sketches/sec (1024^2 ViT-H encode + all-box decode)"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", 1400.0), d.get("hbm_gbs", 6650.0), "measured"
    return 1400.0, 6650.0, "fallback"


def gemm_traffic():
    """DRAM bytes per encoder-GEMM launch from the committed `ncu --set full` capture (tools/ncu_summary.py traffic);
    None if no capture has been summarised."""
    p = os.path.join(ROOT, "profiles", "gemm_traffic.json")
    if not os.path.exists(p):
        return None
    return json.load(open(p)).get("dram_bytes_per_launch")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def synth_inputs(rank: int, n: int):
    import numpy as np

    from inklayer_b200.synth import synthetic_boxes, synthetic_sketch

    imgs = [synthetic_sketch(rank * 1000 + i) for i in range(n)]
    boxes = [synthetic_boxes(rank * 1000 + i, BOXES) for i in range(n)]
    return imgs, boxes, np


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """The reference's own algorithm on the host cores: oracle/sam_oracle.py, the CPU restatement pinned to the
    vendored segment_anything by tests/golden (the reference tree itself is not present on the GPU box).
    Each step = ONE sketch of the workload (ViT-H fp32, 16 boxes), bounded so the run ends within a few minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch

    from inklayer_b200.config import CONFIGS
    from inklayer_b200.weights import synthetic_state_dict
    from oracle import sam_oracle as O

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    imgs, boxes, np = synth_inputs(0, 1)
    pred = O.OraclePredictor(O.CONFIGS[MODEL], synthetic_state_dict(CONFIGS[MODEL], 0))

    def one():
        t0 = time.perf_counter()
        pred.set_image(imgs[0])
        tb = O.apply_boxes(torch.from_numpy(boxes[0]), imgs[0].shape[:2])
        pred.predict_torch(tb)
        return time.perf_counter() - t0

    warm = min(args.warmup, 1)
    t_w = [one() for _ in range(warm)]
    est = t_w[0] if t_w else 20.0
    steps = max(1, min(args.steps, int(150.0 / max(est, 1e-3))))
    ts = [one() for _ in range(steps)]
    v = steps / sum(ts)
    sample = f"{steps} step(s) of 1 sketch each (ViT-H fp32, {BOXES} boxes), {warm} warm-up; torch CPU ops, {cores} threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "sketches/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": 1e3 * sum(ts) / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"SAM {MODEL} 1024x1024 sketch encode + {BOXES}-box decode + post-process", "model": MODEL},
        "cpu_baseline": {"value": v, "unit": "sketches/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "sketches/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------ B200 arm
def cpu_baseline_sample():
    """Oracle (CPU port of the reference path) on the host cores: one ViT-H sketch with 16 boxes, timed once after
    no warm-up (about 10-30 s of CPU work)."""
    import torch

    from inklayer_b200.config import CONFIGS
    from inklayer_b200.weights import synthetic_state_dict
    from oracle import sam_oracle as O

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    imgs, boxes, np = synth_inputs(0, 1)
    pred = O.OraclePredictor(O.CONFIGS[MODEL], synthetic_state_dict(CONFIGS[MODEL], 0))
    t0 = time.perf_counter()
    pred.set_image(imgs[0])
    pred.predict_torch(O.apply_boxes(torch.from_numpy(boxes[0]), imgs[0].shape[:2]))
    dt = time.perf_counter() - t0
    return {"value": 1.0 / dt, "unit": "sketches/s", "cores": cores, "kind": "port",
            "sample": f"1 sketch (ViT-H fp32, {BOXES} boxes) through oracle/sam_oracle.py in {dt:.1f} s, torch CPU ops, {cores} threads"}


def run_b200(args):
    import torch
    import torch.distributed as dist

    from inklayer_b200 import CONFIGS, build_sam_from_state_dict
    from inklayer_b200.pipeline import SketchSegmenter
    from inklayer_b200.weights import synthetic_state_dict

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    cfg = CONFIGS[MODEL]
    sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to(dev)
    imgs, boxes, np = synth_inputs(rank, BATCH)  # sketches shard by image: each rank has its own
    seg = SketchSegmenter(sam, max_batch=BATCH)
    d_imgs = torch.from_numpy(np.stack(imgs)).to(dev)
    d_boxes = [torch.from_numpy(b).to(dev) for b in boxes]
    size = (1024, 1024)

    d_masks = torch.empty(BATCH * BOXES * 1024 * 1024, dtype=torch.uint8, device=dev)
    sizes = [size] * BATCH

    def step_resident():
        feats = sam.encode(d_imgs, graph=True)  # the encoder's ~200 launches replay from a CUDA graph
        return sam.decode_batch(feats, d_boxes, sizes, sizes, masks_out=d_masks)

    def step_e2e(steps=1):
        # the public API: host uint8 sketches + host boxes in, host bool masks out; `steps` batches stream through the
        # segmenter's copy/compute pipeline (H2D from pinned staging and D2H of every mask are inside the timed region)
        n = 0
        for res in seg.segment_stream((imgs, boxes) for _ in range(steps)):
            n += len(res)
        assert n == steps * BATCH

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps, profile=False):
        barrier()
        l0 = sam.launch_count
        if profile:
            sam.profile(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = sam.launch_count - l0
        prof = sam.profile_read() if profile else None
        if profile:
            sam.profile(False)
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), launches, prof

    for _ in range(max(args.warmup, 3)):
        step_resident()
    clocks = ClockSampler(local) if rank == 0 else None
    ms, launches, _ = timed(step_resident, args.steps)
    clk = clocks.stop() if clocks else None
    # the same steps again with the library's per-launch CUDA events on (stream launches): the roofline breakdown
    ms_p, _, prof = timed(step_resident, args.steps, profile=True)
    # encoder alone (config 2 of BASELINE.json: batch 8, bf16 operands) for the tensor-core utilisation figure
    ms_enc, _, _ = timed(lambda: sam.encode(d_imgs, graph=True), args.steps)
    step_e2e(2 * (seg.depth + 1) + 1)  # every pipeline slot has captured its encoder graph before the timed region
    seg.h2d_bytes = seg.d2h_bytes = 0
    ms_e2e, _, _ = timed(lambda: step_e2e(args.steps), 1)

    if rank == 0:
        peak_tf, peak_gbs, peak_src = measured_peaks()
        n_sk = BATCH * world * args.steps
        g_ms, g_flop, g_n = prof["encoder_gemm"]
        # algorithmic HBM bytes of the 4 GEMMs of a ViT block at this batch: A + W + output, + the fp32 residual and the
        # bf16 copy of the stream for proj / lin2 (LayerNorm fold; row statistics are 0.3 % of that)
        M, D = BATCH * 4096, cfg.embed_dim
        blk = ((M * D * 2 + 3 * D * D * 2 + M * 3 * D * 2) + (M * D * 2 + D * D * 2 + 2 * M * D * 4 + M * D * 2) +
               (M * D * 2 + 4 * D * D * 2 + M * 4 * D * 2) + (M * 4 * D * 2 + 4 * D * D * 2 + 2 * M * D * 4 + M * D * 2))
        g_bytes = blk * cfg.depth * args.steps
        a_ms, a_flop, a_n = prof["encoder_attention"]
        d_ms, d_flop, d_n = prof["decoder_gemm"]
        p_ms, p_bytes, p_n = prof["postprocess"]
        achieved = g_flop / (g_ms * 1e-3) / 1e12 if g_ms > 0 else 0.0
        enc_tf = ENC_GFLOP[MODEL] * BATCH * args.steps / (ms_enc * 1e-3) / 1e3
        line = {
            "metric": METRIC, "value": n_sk / (ms * 1e-3), "unit": "sketches/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"SAM {MODEL}: {BATCH} synthetic 1024x1024 sketches per step per GPU, image encoder + "
                                   f"{BOXES}-box mask decode + fused post-process per sketch; random-init weights",
                       "model": MODEL, "global_batch": BATCH * world, "boxes_per_sketch": BOXES, "parallelism": f"image-sharded x{world}, no collective",
                       "l2": "per-step working set (1.2 GB of activations) exceeds the 126 MB L2; no flush needed"},
            "clocks": clk,
            "e2e": {"value": n_sk / (ms_e2e * 1e-3), "unit": "sketches/s", "h2d_bytes_per_step": seg.h2d_bytes // args.steps,
                    "d2h_bytes_per_step": seg.d2h_bytes // args.steps},
            "gpu_launches": launches,
            "roofline": {"bound": "tensor", "kernel": "gemm_tc_kernel (encoder instantiations)", "achieved": achieved, "peak": peak_tf,
                         "unit": "TFLOP/s", "frac": achieved / peak_tf, "traffic": gemm_traffic(),
                         "traffic_unit": "DRAM bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum, profiles/gemm_traffic.json)",
                         "algorithmic_bytes_per_launch": g_bytes / max(g_n, 1), "peak_source": peak_src + " sustained bf16",
                         "launches": g_n, "avg_launch_ms": g_ms / max(g_n, 1), "share_of_step": g_ms / ms_p},
            "encoder": {"ms_per_batch": ms_enc / args.steps, "images_per_s": BATCH * args.steps / (ms_enc * 1e-3),
                        "tflops": enc_tf, "frac_of_peak": enc_tf / peak_tf,
                        "attention_tflops": a_flop / (a_ms * 1e-3) / 1e12 if a_ms > 0 else 0.0, "attention_share_of_step": a_ms / ms_p},
            "decoder": {"gemm_tflops": d_flop / (d_ms * 1e-3) / 1e12 if d_ms > 0 else 0.0, "gemm_share_of_step": d_ms / ms_p,
                        "postprocess_gbs": p_bytes / (p_ms * 1e-3) / 1e9 if p_ms > 0 else 0.0,
                        "postprocess_frac_of_hbm_peak": (p_bytes / (p_ms * 1e-3) / 1e9) / peak_gbs if p_ms > 0 else 0.0},
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_sample()
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the ~20 s oracle timing (profiling runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
