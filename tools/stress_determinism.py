#!/usr/bin/env python
"""Repeats the single-GPU stage on one read set and reports any statistic or checksum (D1 and D2) that differs between
passes: record order inside the bins depends on atomics, the result must not."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mhm2_kmer_analysis_v2_b200 as K  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--reads", type=int, default=4_000_000)
ap.add_argument("--genomes", type=int, default=20)
ap.add_argument("--k", type=int, default=33)
ap.add_argument("--passes", type=int, default=60)
ap.add_argument("--ext", type=int, default=1)
a = ap.parse_args()
reads = K.generate("C2", a.reads, n_genomes=a.genomes)
st = K.KcountStage(a.k, keep_ext_counts=bool(a.ext), est_kmers=reads.counted_kmers(a.k) + 2 * len(reads))
st.add_reads(reads)
seen = {}
skip = {"ms_parse", "ms_insert", "ms_finalize", "kernel_launches", "ms_parse_kernel", "ms_insert_kernel", "ms_table_clear", "ms_refine_kernel",
        "ms_exchange_copy", "result_bytes", "insert_kernel_launches"}
for i in range(a.passes):
    st.reset()
    s = st.run()
    s = {k: v for k, v in s.items() if k not in skip}
    s["checksum_d1"] = st.checksum(False)
    if a.ext:
        s["checksum_d2"] = st.checksum(True)
    for k, v in s.items():
        seen.setdefault(k, {}).setdefault(v, []).append(i)
bad = {k: {str(v): p[:8] for v, p in d.items()} for k, d in seen.items() if len(d) > 1}
print("passes", a.passes, "differing:", bad if bad else "none")
print({k: list(d)[0] for k, d in seen.items() if len(d) == 1})
