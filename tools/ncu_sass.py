#!/usr/bin/env python
"""Per-SASS-instruction execution counts of one kernel of an .ncu-rep: prints instructions matching a regex, and the
total warp instructions between consecutive matches (a cheap way to size loop bodies).
usage: python tools/ncu_sass.py rep.ncu-rep kernel-regex [opcode-regex]"""
import csv
import io
import re
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
pat = re.compile(sys.argv[3] if len(sys.argv) > 3 else "ATOMS|VOTE|BAR")
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = {x: i for i, x in enumerate(rows[1])}
tot = 0
acc = 0
for r in rows[2:]:
    if len(r) < 10:
        continue
    n = int(r[hdr["Instructions Executed"]] or 0)
    t = int(r[hdr["Thread Instructions Executed"]] or 0)
    tot += n
    acc += n
    if pat.search(r[1]):
        print(f"{r[0][-5:]} {n/1e6:9.2f}M thr={t/max(n,1):5.1f}  since last match {acc/1e6:9.1f}M   {r[1].strip()[:80]}")
        acc = 0
print(f"total {tot/1e9:.3f} G warp instructions")
