"""GPU, config scale: the CUDA stage against the REFERENCE's own CPU kcount on BASELINE's configs -- C1 in full (1 M reads,
k = 21) and 2 M-read prefixes of C2 (k = 21/33/55), C3 (paired, k = 21/55), C4 (k = 77/99, m = 15 and m = 27) and C5 (1 %
error, k = 21/33), plus 10-genome variants at the full workloads' coverage ("parity on C1 in full and on
deterministic subsamples of C2-C5").

tests/golden/config_checksums.json holds what oracle/_ref/kcount_ref (the reference's sources, compiled unmodified)
computed for every case: table entries, entries with count >= 2, survivors, sum(count), max count and the
order-independent checksums of dumps D1 (survivors: k-mer, count, left, right) and D2 (every entry with count >= 2 with
its 8 extension counters and chosen extensions incl. X / F); tests/golden/make_config_checksums.py wrote it. The reads are
a seeded stream regenerated here. Every table variant of the stage -- bin-resident shared-memory tables (default), one HBM
table, HBM table behind the exact singleton filter -- must reproduce all of it bit for bit through the C ABI.
Reference behaviour pinned: src/kcount/kcount_cpu.cpp:309-355 (insert), :524-601 (purge / extensions),
src/kcount/kmer_dht.cpp:117-119, 193-196 (m rule, owner)."""
import json
import os

import numpy as np
import pytest

import mhm2_kmer_analysis_v2_b200 as K

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "config_checksums.json")) as f:
    CASES = json.load(f)["cases"]

_reads_cache = {}


def _reads(case):
    key = (case["config"], json.dumps(case["synth"], sort_keys=True), case["first_read"], case["n_reads"])
    if key not in _reads_cache:
        _reads_cache.clear()  # one read set at a time (300 MB each)
        _reads_cache[key] = K.generate(None, case["n_reads"], first=case["first_read"], **case["synth"])
    return _reads_cache[key]


def _variants(case):
    out = [("resident", {})]
    if case["config"] == "C4":
        out += [("resident_m15", dict(m=15)), ("resident_m27", dict(m=27))]
    out += [("hbm_table", dict(global_table=True)), ("filter", dict(use_filter=True))]
    if case["config"] == "C4":
        out += [("filter_m15", dict(use_filter=True, m=15))]
    return out


@pytest.mark.parametrize("case", CASES, ids=[f"{c['case']}_k{c['k']}" for c in CASES])
def test_config_scale_equals_reference(case):
    reads = _reads(case)
    k = case["k"]
    assert reads.counted_kmers(k) == case["counted_kmers"]
    raw = int(np.maximum(reads.lens.astype(np.int64) - k + 1, 0).sum())
    for name, kw in _variants(case):
        st = K.KcountStage(k, dmin_thres=case["dmin"], keep_ext_counts=True, est_kmers=raw, **kw)
        try:
            st.add_reads(reads)
            s = st.run()
            got = dict(entries=s["unique_kmers"] + s["filtered_singletons"], ge2=s["ge2_kmers"], survivors=s["final_kmers"], sum_count=s["sum_counts"],
                       max_count=s["max_count"], checksum_d1=st.checksum(False), checksum_d2=st.checksum(True), inserted=s["inserted_kmers"],
                       dropped=s["dropped"])
            want = dict(entries=case["entries"], ge2=case["ge2"], survivors=case["survivors"], sum_count=case["sum_count"],
                        max_count=case["max_count"], checksum_d1=case["checksum_d1"], checksum_d2=case["checksum_d2"],
                        inserted=case["counted_kmers"], dropped=0)
            if kw.get("use_filter"):
                # the filter keeps true singletons (and nothing else) out of the table: the table population is not comparable,
                # everything downstream of the purge is
                got.pop("entries"), want.pop("entries")
            assert got == want, f"{case['case']} k={k} variant {name}: {got} != reference {want}"
            if name == "resident":
                assert s["local_bins"] >= 65536, "config-scale input must take the two-level partition (parse -> refine -> count)"
        finally:
            st.close()


def test_repeated_passes_give_one_result():
    """Record order inside the bins depends on atomics; nothing the stage reports may (split_passes aside): 20 passes over
    one read set, every statistic and both dump checksums identical."""
    case = next(c for c in CASES if c["case"] == "C2_10genomes_2M")
    reads = _reads(case)
    k = case["k"]
    st = K.KcountStage(k, keep_ext_counts=True, est_kmers=reads.counted_kmers(k) + 2 * len(reads))
    st.add_reads(reads)
    seen = set()
    names = ("inserted_kmers", "unique_kmers", "ge2_kmers", "final_kmers", "purged_kmers", "sum_counts", "max_count", "supermers_in")
    for _ in range(20):
        st.reset()
        s = st.run()
        seen.add(tuple(s[n] for n in names) + (st.checksum(False), st.checksum(True)))
    st.close()
    assert len(seen) == 1, seen
    assert next(iter(seen))[-2:] == (case["checksum_d1"], case["checksum_d2"])
