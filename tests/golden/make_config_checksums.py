#!/usr/bin/env python
"""Config-scale pins: runs the REFERENCE's own CPU kcount (oracle/_ref/kcount_ref, the reference sources compiled
unmodified behind oracle/ref_shim) on BASELINE's configs -- C1 in full, deterministic 2 M-read prefixes of C2-C5 (SURVEY.md
8d) -- and commits what it computed as tests/golden/config_checksums.json: table entries, entries with count >= 2,
survivors, sum(count), max count and the order-independent checksums of the dumps D1 / D2 (kco_entry_mix summed, the
function kc_result_checksum evaluates on the device). The reads are NOT stored: they are a seeded stream
(libkc_synth.so; read i depends only on (seed, i)), every entry names its generator options, and tests/test_gpu_configs.py
regenerates them on the GPU box. Only runs where /root/reference exists:

    make -C oracle ref && python tests/golden/make_config_checksums.py [case ...]

The result of the stage does not depend on the minimizer length or the number of owners, so one reference
run pins every (m, n_ranks, table variant) combination of a (reads, k, dmin) case.
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import mhm2_kmer_analysis_v2_b200 as K  # noqa: E402
import oracle as O  # noqa: E402

OUT = os.path.join(HERE, "config_checksums.json")
TMP = "/dev/shm/kc_cfg" if os.path.isdir("/dev/shm") else "/tmp/kc_cfg"

# name -> (config of reads.CONFIGS, synth overrides, reads, k list, dmin)
CASES = {
    "C1_full": ("C1", {}, 1_000_000, (21,), 2),
    "C2_prefix2M": ("C2", {}, 2_000_000, (21, 33, 55), 2),
    # the same community at the full workload's coverage (10x): a prefix of a 100-genome mix is ~1x deep, this one
    # exercises counts around the dynamic-min-depth edges and the extension rule at depth
    "C2_10genomes_2M": ("C2", {"n_genomes": 10}, 2_000_000, (33,), 2),
    "C3_prefix2M": ("C3", {}, 2_000_000, (21, 55), 2),
    "C4_prefix2M": ("C4", {}, 2_000_000, (77, 99), 2),
    "C4_10genomes_2M": ("C4", {"n_genomes": 10}, 2_000_000, (77, 99), 2),
    "C5_prefix2M": ("C5", {}, 2_000_000, (21,), 2),
    "C5_10genomes_2M": ("C5", {"n_genomes": 10}, 2_000_000, (21, 33), 2),
}


def run_case(name, ranks):
    cfg, over, n, ks, dmin = CASES[name]
    os.makedirs(TMP, exist_ok=True)
    reads = K.generate(cfg, n, **over)
    rp = os.path.join(TMP, name + ".reads")
    reads.save(rp)
    out = []
    for k in ks:
        t0 = time.time()
        pre = os.path.join(TMP, f"{name}_k{k}")
        s = O.run_reference(rp, k, pre, extended=True, dmin=dmin, ranks=ranks)
        assert s["dropped"] in (0, -1), s
        assert s["occurrences"] == reads.counted_kmers(k)
        d1, d2 = O.load_dump(pre + ".d1"), O.load_dump(pre + ".d2")
        rec = {
            "case": name, "config": cfg, "synth": dict(K.CONFIGS[cfg]["synth"], **over), "first_read": 0, "n_reads": n, "k": k, "dmin": dmin,
            "counted_kmers": int(s["occurrences"]), "entries": int(s["entries"]), "ge2": len(d2), "survivors": len(d1),
            "sum_count": int(d1.counts.astype("u8").sum()), "max_count": int(s["max_count"]),
            "checksum_d1": O.result_checksum(d1, False), "checksum_d2": O.result_checksum(d2, True),
            "reference_ranks": ranks,
        }
        for suf in (".d1", ".d2"):
            os.unlink(pre + suf)
        print(json.dumps(rec), f"# {time.time() - t0:.0f} s", flush=True)
        out.append(rec)
    os.unlink(rp)
    return out


if __name__ == "__main__":
    if not os.path.exists(O.REF_BIN):
        raise SystemExit("oracle/_ref/kcount_ref missing: run `make -C oracle ref` where /root/reference exists")
    names = sys.argv[1:] or list(CASES)
    ranks = min(len(os.sched_getaffinity(0)), 8)
    have = {}
    if os.path.exists(OUT):
        with open(OUT) as f:
            for r in json.load(f)["cases"]:
                have[(r["case"], r["k"])] = r
    for nm in names:
        for r in run_case(nm, ranks):
            have[(r["case"], r["k"])] = r
        with open(OUT, "w") as f:
            json.dump({"generator": "tests/golden/make_config_checksums.py (oracle/_ref/kcount_ref = reference sources, unmodified)",
                       "cases": [have[key] for key in sorted(have)]}, f, indent=1)
