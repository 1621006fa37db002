"""Prints the table of tools/decode_kernels.sh from its ncu csv."""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ik, im, iv, iid, iu = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID"), hdr.index("Metric Unit")
d = {}
for r in rows[1:]:
    v = float(r[iv].replace(",", ""))
    u = r[iu]
    if u in ("ns", "nsecond"): v /= 1e3
    if u in ("ms", "msecond"): v *= 1e3
    if u == "byte": v /= 1e6
    if u == "Kbyte": v /= 1e3
    if u == "Gbyte": v *= 1e3
    d.setdefault((int(r[iid]), r[ik].replace("inksam::", "").replace("void ", "")[:44]), {})[r[im]] = v
tot = 0.0
for (i, k), m in sorted(d.items()):
    t = m["gpu__time_duration.sum"]
    tot += t
    print(f"{k:44s} {t:8.1f} us  rd {m['dram__bytes_read.sum']:7.1f} MB  wr {m['dram__bytes_write.sum']:7.1f} MB  "
          f"l1 {m['l1tex__throughput.avg.pct_of_peak_sustained_elapsed']:5.1f}  tensor {m['sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active']:5.1f}  "
          f"issue {m['smsp__issue_active.avg.pct_of_peak_sustained_active']:5.1f}")
print(f"sum {tot:.1f} us")
