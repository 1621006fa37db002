#!/usr/bin/env python
"""Warp-stall summary of one kernel of an `ncu --set full --import-source on` report (source page, SASS view):
   tools/ncu_stalls.py report.ncu-rep [top_n]
Prints the stall-reason totals, the samples between synchronisation instructions (mbarrier try_wait / arrive, bar.sync,
tcgen05.commit, exit: the 'phases' of a warp-specialised kernel) and the hottest instructions."""
import csv
import io
import subprocess
import sys


def main():
    path = sys.argv[1]
    top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    # one section per profiled launch: a "Kernel Name" row, a header row, then one row per SASS instruction
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    seen = set()
    for k, i0 in enumerate(starts):
        i1 = starts[k + 1] if k + 1 < len(starts) else len(rows)
        key = (rows[i0][1], tuple(tuple(r[:4]) for r in rows[i0 + 2:i0 + 6]))  # (ncu lists every launch twice)
        if key in seen:
            continue
        seen.add(key)
        section(rows[i0:i1], top_n)
        print()


def section(rows, top_n):
    print("#", rows[0][1][:150])
    hdr, data = rows[1], [r for r in rows[2:] if len(r) > 5]
    isamp, ia, iex = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    base = int(data[0][0], 16)
    tot = sum(int(r[isamp]) for r in data)
    agg = {s: sum(int(r[hdr.index(s)]) for r in data if r[hdr.index(s)].isdigit()) for s in stalls}
    print(f"samples {tot}, SASS instructions {len(data)}")
    print("stall reasons: " + ", ".join(f"{s[6:]} {100 * v / tot:.1f}%" for s, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    print("-- samples between synchronisation instructions")
    seg, seg_start = 0, 0
    for r in data:
        off, s, src = int(r[0], 16) - base, int(r[isamp]), r[ia].strip()
        seg += s
        if any(m in src for m in ("SYNCS.PHASECHK", "BAR.SYNC", "EXIT", "SYNCS.ARRIVE", "UTCBAR")):
            if seg >= 0.01 * tot:
                print(f"{seg_start:05x}-{off:05x} {seg:6d} {100 * seg / tot:5.1f}%  ends at: {src[:72]}  (executed {r[iex]})")
            seg, seg_start = 0, off + 16
    print("-- hottest instructions")
    for r in sorted(data, key=lambda r: -int(r[isamp]))[:top_n]:
        st = sorted(((s[6:], int(r[hdr.index(s)])) for s in stalls if r[hdr.index(s)].isdigit() and int(r[hdr.index(s)]) > 0), key=lambda x: -x[1])
        print(f"{int(r[0], 16) - base:05x} {int(r[isamp]):6d} {100 * int(r[isamp]) / tot:5.1f}%  ex={r[iex]:>8s}  {r[ia].strip()[:64]:64s} {st[:2]}")


if __name__ == "__main__":
    main()
