#!/usr/bin/env python
"""Summaries of ncu output for profiles/.

    tools/ncu_summary.py launches gpurun_out/launches.csv            # per-kernel totals of a launch list
    tools/ncu_summary.py full gpurun_out/prof_gemm.ncu-rep [...]     # key metrics of every launch in --set full reports
    tools/ncu_summary.py traffic gpurun_out/prof_gemm.ncu-rep out.json  # mean DRAM bytes per launch (bench.py roofline.traffic)
    tools/ncu_summary.py decoder-traffic gpurun_out/dec_traffic.csv out.json  # DRAM bytes of one decode pass (bench.py decoder.hbm_gbs)
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "launch__registers_per_thread",
    "launch__grid_size",
    "launch__block_size",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
]


def short(name: str) -> str:
    return re.sub(r"\(.*", "", name).replace("void ", "").replace("inksam::", "")


def launches(path: str) -> None:
    lines = [l for l in open(path) if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ix = {h: i for i, h in enumerate(hdr)}
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for row in r:
        if len(row) < len(hdr):
            continue
        v = float(row[ix["Metric Value"]].replace(",", ""))
        u = row[ix["Metric Unit"]]
        v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
        k = short(row[ix["Kernel Name"]])
        agg[k][0] += 1
        agg[k][1] += v
        tot += v
    print(f"# {path}: {sum(a[0] for a in agg.values())} launches, {tot / 1e3:.3f} ms total (ncu per-launch times: cold cache, serialised)")
    print(f"{'total ms':>10} {'count':>6} {'avg us':>9} {'share':>7}  kernel")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{t / 1e3:10.3f} {n:6d} {t / n:9.1f} {100 * t / tot:6.1f}%  {k}")


def full(paths) -> None:
    for p in paths:
        out = subprocess.run(["ncu", "-i", p, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        hdr, units = rows[0], rows[1]
        ix = [hdr.index(k) for k in KEYS if k in hdr]
        print(f"# {p}")
        print(" | ".join(["kernel"] + [f"{hdr[i]} [{units[i]}]" for i in ix]))
        for row in rows[2:]:
            print(" | ".join([short(row[hdr.index("Kernel Name")])] + [row[i] for i in ix]))
        print()


def traffic(path: str, out_json: str) -> None:
    """Mean of dram__bytes_read.sum + dram__bytes_write.sum over the bf16 encoder GEMM launches of a --set full capture
    (a whole number of layers: qkv, proj, lin1, lin2 repeating)."""
    import json

    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = []
    for row in rows[2:]:
        name = short(row[ik])
        if not re.match(r"gemm_tc_kernel<\d, 0,", name):   # bf16 instantiations = the encoder
            continue
        per.append({"kernel": name, "bytes": float(row[ir]) * scale[units[ir]] + float(row[iw]) * scale[units[iw]]})
    mean = sum(p["bytes"] for p in per) / max(len(per), 1)
    json.dump({"source": path, "launches": len(per), "dram_bytes_per_launch": mean, "per_launch": per}, open(out_json, "w"), indent=1)
    print(f"{len(per)} launches, mean DRAM traffic {mean / 1e6:.1f} MB per launch -> {out_json}")


def decoder_traffic(path: str, out_json: str) -> None:
    """Sum of dram__bytes_read.sum + dram__bytes_write.sum (and of gpu__time_duration.sum) over every launch of ONE decode
    pass: `ncu --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --profile-from-start off`
    around tools/decode_pass.py.  bench.py divides the byte total by the pass's live CUDA-event time (decoder.hbm_gbs)."""
    import json

    lines = [l for l in open(path) if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ix = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0,
             "ms": 1e3, "msecond": 1e3}
    per = collections.OrderedDict()
    for row in r:
        if len(row) < len(hdr):
            continue
        key = (row[ix["ID"]], short(row[ix["Kernel Name"]]))
        d = per.setdefault(key, {"bytes": 0.0, "us": 0.0})
        v = float(row[ix["Metric Value"]].replace(",", "")) * scale.get(row[ix["Metric Unit"]], 1.0)
        if row[ix["Metric Name"]].startswith("dram__bytes"):
            d["bytes"] += v
        elif row[ix["Metric Name"]].startswith("gpu__time_duration"):
            d["us"] += v
    agg = collections.OrderedDict()
    for (_, k), d in per.items():
        a = agg.setdefault(k, {"kernel": k, "launches": 0, "bytes": 0.0, "us": 0.0})
        a["launches"] += 1; a["bytes"] += d["bytes"]; a["us"] += d["us"]
    total = sum(d["bytes"] for d in per.values())
    json.dump({"source": path, "launches": len(per), "dram_bytes_per_pass": total, "ncu_us_per_pass": sum(d["us"] for d in per.values()),
               "per_kernel": sorted(agg.values(), key=lambda a: -a["bytes"])}, open(out_json, "w"), indent=1)
    print(f"{len(per)} launches, DRAM traffic {total / 1e6:.1f} MB per decode pass -> {out_json}")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    elif sys.argv[1] == "traffic":
        traffic(sys.argv[2], sys.argv[3])
    elif sys.argv[1] == "decoder-traffic":
        decoder_traffic(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[2:])
