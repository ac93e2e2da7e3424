#!/usr/bin/env python
"""Per CUDA source line view of an .ncu-rep (needs -lineinfo and --import-source on): samples, warp instructions, threads/inst.
usage: python tools/ncu_lines.py rep.ncu-rep [n_top] [file-substring]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
n_top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
want = sys.argv[3] if len(sys.argv) > 3 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
fname, hdr, rows = "", None, []
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] in ("File Path", "File Name"):
        fname = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = {x: i for i, x in enumerate(r)}
    elif hdr and len(r) > 10 and r[2] == "-" and r[0].isdigit():
        rows.append((fname, int(r[0]), r[1].strip(), int(r[hdr["# Samples"]] or 0), int(r[hdr["Instructions Executed"]] or 0),
                     int(r[hdr["Thread Instructions Executed"]] or 0), {k[6:]: int(r[i] or 0) for k, i in hdr.items() if k.startswith("stall_") and "Not Issued" not in k}))
rows = [r for r in rows if want in r[0]]
ts = sum(r[3] for r in rows) or 1
ti = sum(r[4] for r in rows) or 1
print(f"total samples {ts}, warp instructions {ti/1e9:.2f} G")
for r in sorted(rows, key=lambda r: -r[3])[:n_top]:
    st = sorted(r[6].items(), key=lambda kv: -kv[1])[:2]
    thr = r[5] / r[4] if r[4] else 0
    print(f"{r[3]/ts*100:5.1f}% smp {r[4]/ti*100:5.1f}% inst thr={thr:4.1f} {r[0]}:{r[1]:<4d} {r[2][:70]:70s} {st[0][0]}:{st[0][1]} {st[1][0]}:{st[1][1]}")
# optional: aggregate by line ranges of one file: --ranges file:lo-hi,lo-hi,...
for a in sys.argv[4:]:
    if a.startswith("--ranges="):
        f, rs = a[9:].split(":")
        for r in rs.split(","):
            lo, hi = map(int, r.split("-"))
            sel = [x for x in rows if x[0] == f and lo <= x[1] <= hi]
            print(f"{f}:{lo}-{hi}: {sum(x[3] for x in sel)/ts*100:5.1f}% samples {sum(x[4] for x in sel)/ti*100:5.1f}% inst "
                  f"thr={sum(x[5] for x in sel)/max(sum(x[4] for x in sel),1):.1f}")
        oth = [x for x in rows if x[0] != f]
        print(f"other files: {sum(x[3] for x in oth)/ts*100:5.1f}% samples {sum(x[4] for x in oth)/ti*100:5.1f}% inst")
        byf = {}
        for x in oth:
            byf.setdefault(x[0], [0, 0])
            byf[x[0]][0] += x[3]
            byf[x[0]][1] += x[4]
        for k2, v in sorted(byf.items(), key=lambda kv: -kv[1][1]):
            print(f"   {k2}: {v[0]/ts*100:5.1f}% samples {v[1]/ti*100:5.1f}% inst")
