#!/usr/bin/env python
"""Stall-sample hot spots of an .ncu-rep captured with --import-source on:
python scripts/ncu_hotspots.py file.ncu-rep [N]  -> stall reasons over the kernel + the N hottest SASS instructions."""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 12
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
print(rows[0][1][:160])
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = [r for r in rows[2:] if r[idx["# Samples"]].isdigit()]
S = idx["# Samples"]
total = sum(int(r[S]) for r in data)
tot = {c: sum(int(r[idx[c]] or 0) for r in data) for c in cols}
T = sum(tot.values()) or 1
print(f"{len(data)} SASS instructions, {total} samples")
print("stall reasons: " + ", ".join(f"{c[6:]} {100 * v / T:.1f}%" for c, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v * 200 > T))
for i in sorted(range(len(data)), key=lambda i: -int(data[i][S]))[:top]:
    r = data[i]
    why = ", ".join(f"{c[6:]} {r[idx[c]]}" for c in cols if int(r[idx[c]] or 0) * 10 > int(r[S]))
    print(f"  #{i:5d} {100 * int(r[S]) / total:5.1f}%  {r[idx['Source']].strip()[:70]:70s} [{why}]")
