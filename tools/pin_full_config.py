#!/usr/bin/env python
"""Full-size pin of a BASELINE config: runs the REFERENCE's CPU kcount (oracle/_ref/kcount_ref) on all reads of the
workload on a host with enough memory (the GPU box: the reference's table for C2 needs ~60 GB) and writes what it computed
-- entries, survivors, sum(count), max count, D1 checksum -- as one JSON record. bench.py compares its headline result with
tests/golden/full_config_pins.json when the configuration matches. Test infrastructure; no product code is loaded."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np  # noqa: E402

import oracle as O  # noqa: E402
from mhm2_kmer_analysis_v2_b200 import reads as R  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="C2")
ap.add_argument("--k", type=int, default=33)
ap.add_argument("--reads", type=int, default=20_000_000)
ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "full_config_pins.json"))
a = ap.parse_args()
cores = min(len(os.sched_getaffinity(0)), 64)
t0 = time.time()
reads = R.generate(a.workload, a.reads)
pre = f"/dev/shm/kc_pin_{os.getpid()}"
reads.save(pre + ".reads")
occ = reads.counted_kmers(a.k)
del reads
s = O.run_reference(pre + ".reads", a.k, pre, ranks=cores)
os.unlink(pre + ".reads")
assert s["dropped"] in (0, -1) and s["occurrences"] == occ, s
d1 = O.load_dump(pre + ".d1")
os.unlink(pre + ".d1")
rec = {"workload": a.workload, "k": a.k, "first_read": 0, "n_reads": a.reads, "synth": R.CONFIGS[a.workload]["synth"], "counted_kmers": occ,
       "entries": int(s["entries"]), "survivors": len(d1), "sum_count": int(d1.counts.astype(np.uint64).sum()),
       "max_count": int(s["max_count"]), "checksum_d1": f"{O.result_checksum(d1, False):016x}", "reference_ranks": cores,
       "reference_seconds": round(s["t_dht_s"] + s["t_analyze_s"], 2), "wall_seconds": round(time.time() - t0, 1),
       "generator": "tools/pin_full_config.py (oracle/_ref/kcount_ref = reference sources, unmodified)"}
have = []
if os.path.exists(a.out):
    with open(a.out) as f:
        have = json.load(f)["pins"]
have = [h for h in have if (h["workload"], h["k"], h["n_reads"]) != (a.workload, a.k, a.reads)] + [rec]
os.makedirs(os.path.dirname(a.out), exist_ok=True)
with open(a.out, "w") as f:
    json.dump({"pins": have}, f, indent=1)
print(json.dumps(rec))
