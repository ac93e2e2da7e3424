"""CPU: the oracle (oracle/kcount_oracle.c, a restatement of the reference's CPU kcount) against the golden vectors in
tests/golden/*.npz, which are outputs of the REFERENCE's own sources compiled here (tests/golden/make_golden.py ->
oracle/_ref/kcount_ref). The reference ships no tests for this path, so these files are the pin.
"""
import glob
import json
import os

import numpy as np
import pytest

import mhm2_kmer_analysis_v2_b200 as K

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def load_golden(path):
    d = np.load(path)
    k, dmin = int(d["k"]), int(d["dmin"])
    if "recipe" in d.files:
        r = json.loads(str(d["recipe"]))
        assert r["recipe"] == "saturation"
        one = K.ReadSet.from_strings([r["seq"]])
        n = r["copies"]
        reads = K.ReadSet(np.tile(one.bytes, n), np.arange(n, dtype=np.uint64) * np.uint64(len(r["seq"])),
                          np.full(n, len(r["seq"]), np.uint16))
    else:
        reads = K.ReadSet(d["bytes"], d["offsets"], d["lens"])
    return d, k, dmin, reads, json.loads(str(d["summary"]))


def check_against_golden(res_d1, res_d2, d, what):
    for name, res, pre in (("D1", res_d1, "d1_"), ("D2", res_d2, "d2_")):
        assert len(res) == len(d[pre + "counts"]), f"{what} {name}: {len(res)} entries vs reference {len(d[pre + 'counts'])}"
        assert np.array_equal(res.keys, d[pre + "keys"]), f"{what} {name}: keys differ"
        assert np.array_equal(res.counts, d[pre + "counts"]), f"{what} {name}: counts differ"
        assert np.array_equal(res.left.view(np.uint8), d[pre + "left"]), f"{what} {name}: left ext differs"
        assert np.array_equal(res.right.view(np.uint8), d[pre + "right"]), f"{what} {name}: right ext differs"
    assert np.array_equal(res_d2.ext8, d["d2_ext8"]), f"{what}: extension counters differ"


def test_golden_files_present():
    names = {os.path.basename(p) for p in GOLDEN}
    for k in (21, 33, 55, 77, 99):
        assert f"synth_k{k}.npz" in names and f"edge_k{k}.npz" in names
    assert "saturation_k21.npz" in names


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_equals_reference_output(oracle_mod, path):
    d, k, dmin, reads, summary = load_golden(path)
    o = oracle_mod.Oracle(k, dmin_thres=dmin, slots_per_rank=max(4 * reads.counted_kmers(k), 1 << 12))
    o.add_reads(reads.bytes, reads.offsets, reads.lens)
    o.finalize()
    check_against_golden(o.result(False), o.result(True), d, os.path.basename(path))
    assert o.counted_kmers() == summary["occurrences"] == reads.counted_kmers(k)
    assert o.num_entries() == summary["entries"]
    assert len(o.result(False)) == summary["final_kmers"]
    assert int(o.result(False).counts.astype(np.int64).sum()) == summary["sum_count"]
    assert abs(o.avg_kmer_count() - summary["avg_kmer_count"]) < 1e-5  # the reference prints 6 decimals
    # the checksum both sides of the GPU parity tests use agrees with its numpy restatement over the reference dump
    ref_d2 = oracle_mod.Result(d["d2_keys"], d["d2_counts"], d["d2_left"].view("S1"), d["d2_right"].view("S1"), d["d2_ext8"])
    assert o.checksum(True) == oracle_mod.result_checksum(ref_d2, True)


@pytest.mark.parametrize("name,G", [("synth_k21", 2), ("synth_k33", 3), ("synth_k55", 4), ("synth_k77", 8), ("edge_k99", 5),
                                    ("edge_k22_palindromes", 3)])
def test_owner_partition_is_invisible_in_the_union(oracle_mod, name, G):
    """The table does not depend on the number of owners: G emulated ranks, union == reference 1-rank dump."""
    d, k, dmin, reads, _ = load_golden(os.path.join(os.path.dirname(__file__), "golden", name + ".npz"))
    o = oracle_mod.Oracle(k, dmin_thres=dmin, n_ranks=G, slots_per_rank=max(4 * reads.counted_kmers(k), 1 << 12))
    o.add_reads(reads.bytes, reads.offsets, reads.lens)
    o.finalize()
    check_against_golden(o.result(False), o.result(True), d, f"{name} G={G}")
    per_rank = [o.result(True, rank=r) for r in range(G)]
    assert sum(len(r) for r in per_rank) == len(d["d2_counts"])
    # every key sits on the rank get_kmer_target_rank names
    L = oracle_mod.lib()
    for r, res in enumerate(per_rank):
        for row in res.keys[:50]:
            kw = np.ascontiguousarray(row)
            assert L.kco_owner(kw.ctypes.data, k, o.m, G) == r


def test_supermers_cover_every_interior_kmer_once(oracle_mod):
    """process_seq (kcount_cpu.cpp:74-103): supermers of a read tile the k-mers 1 .. len-k-1, in order, and consecutive
    supermers have different targets."""
    reads = K.generate(None, 200, seed=5, n_genomes=1, genome_len_min=5000, genome_len_max=5000, read_len_min=30, n_rate=0.01)
    for k in (21, 33, 55):
        o = oracle_mod.Oracle(k, n_ranks=7)
        for i in range(len(reads)):
            b = reads.bytes[int(reads.offsets[i]): int(reads.offsets[i]) + int(reads.lens[i])]
            sm = o.read_supermers(b)
            L = len(b)
            if L < k + 2:
                assert sm == []
                continue
            at = 1
            for j, (t, f, n) in enumerate(sm):
                assert f == at and n >= 1 and 0 <= t < 7
                if j:
                    assert sm[j - 1][0] != t
                at += n
            assert at == L - k


def test_get_ext_rule_exhaustive(oracle_mod):
    """ExtCounts::get_ext (kcount_cpu.cpp:173-182): (int)((1.0 - 0.9) * count) == (count - 1) / 10 for every u16 count
    (SURVEY.md section 8), and the X / F / base decisions around the threshold."""
    L = oracle_mod.lib()

    def ext(c4, count, dmin=2):
        a = np.array(c4, np.uint16)
        return L.kco_get_ext(a.ctypes.data, count, dmin, 0.9)

    for count in list(range(0, 300)) + [999, 1000, 1001, 9999, 10000, 10001, 65534, 65535]:
        dyn = max((count - 1) // 10 if count else 0, 2)
        assert dyn == max(int((1.0 - 0.9) * count), 2)
        assert ext([dyn, 0, 0, 0], count) == b"A"
        assert ext([dyn - 1, 0, 0, 0], count) == b"X"
        assert ext([0, 0, dyn, dyn - 1], count) == b"G"
        assert ext([0, dyn, 0, dyn], count) == b"F"
    assert ext([5, 5, 0, 0], 10) == b"F"
    assert ext([0, 0, 0, 0], 2) == b"X"
    assert ext([1, 0, 0, 3], 4) == b"T"
    assert ext([0, 3, 0, 0], 3, dmin=4) == b"X"


def test_kmer_arithmetic_properties(oracle_mod):
    """pack / revcomp / canonical order / minimizer (src/kmer.cpp:303-325, 491-510, 271-277, 350-398)."""
    L = oracle_mod.lib()
    rng = np.random.default_rng(1)
    for k in (21, 31, 33, 55, 63, 77, 99, 127):
        N = (k + 31) // 32
        for _ in range(20):
            s = "".join(rng.choice(list("ACGT"), k))
            w, r, rr = np.zeros(N, np.uint64), np.zeros(N, np.uint64), np.zeros(N, np.uint64)
            L.kco_pack_kmer(s.encode(), k, w.ctypes.data)
            assert np.array_equal(w, K.kcount.pack_kmer(s))
            L.kco_revcomp(w.ctypes.data, k, r.ctypes.data)
            assert np.array_equal(r, K.kcount.pack_kmer(K.kcount.revcomp_kmer_string(s)))
            L.kco_revcomp(r.ctypes.data, k, rr.ctypes.data)
            assert np.array_equal(rr, w)
            # a k-mer and its reverse complement share the minimizer, hence the owner
            m = 15 if k < 40 else 27
            assert L.kco_minimizer(w.ctypes.data, k, m) == L.kco_minimizer(r.ctypes.data, k, m)
            assert L.kco_owner(w.ctypes.data, k, m, 8) == L.kco_owner(r.ctypes.data, k, m, 8)
    # N packs as G (kmer.cpp:173)
    a, b = np.zeros(1, np.uint64), np.zeros(1, np.uint64)
    L.kco_pack_kmer(b"ACNNT", 5, a.ctypes.data)
    L.kco_pack_kmer(b"ACGGT", 5, b.ctypes.data)
    assert a[0] == b[0]
