"""GPU: the CUDA stage (through the C ABI) against the committed golden vectors -- outputs of the reference's own CPU
kcount (tests/golden/make_golden.py). Covers the edge cases the fixtures were built around: reads shorter than k + 2,
N inside k-mers and as extensions, low-quality extensions, palindromes (even k), homopolymers, reverse strands,
saturation of count and extension counters at 65535, dmin_thres 3, the dynamic min-depth threshold at high depth."""
import os

import numpy as np
import pytest

import mhm2_kmer_analysis_v2_b200 as K
from test_oracle_golden import GOLDEN, check_against_golden, load_golden

pytestmark = pytest.mark.gpu
IDS = [os.path.basename(p)[:-4] for p in GOLDEN]


def run_stage(reads, k, dmin, **kw):
    st = K.KcountStage(k, dmin_thres=dmin, keep_ext_counts=True, **kw)
    st.add_reads(reads)
    stats = st.run()
    return st, stats


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_stage_equals_reference_output(path):
    d, k, dmin, reads, summary = load_golden(path)
    st, stats = run_stage(reads, k, dmin)
    check_against_golden(st.result(False), st.result(True), d, os.path.basename(path))
    assert stats["counted_kmers"] == summary["occurrences"]
    assert stats["unique_kmers"] == summary["entries"]
    assert stats["final_kmers"] == summary["final_kmers"]
    assert stats["sum_counts"] == summary["sum_count"]
    assert stats["max_count"] == summary["max_count"]
    assert abs(stats["avg_kmer_count"] - summary["avg_kmer_count"]) < 1e-5
    assert stats["dropped"] == 0
    st.close()


@pytest.mark.parametrize("path", [p for p in GOLDEN if "edge" in p or "saturation" in p or "deep" in p],
                         ids=[i for i in IDS if "edge" in i or "saturation" in i or "deep" in i])
@pytest.mark.parametrize("variant", ["bins", "filter", "global_table"])
def test_variants_equal_reference_output(path, variant):
    d, k, dmin, reads, summary = load_golden(path)
    kw = dict(bins=dict(local_bins=64), filter=dict(use_filter=True), global_table=dict(global_table=True))[variant]
    st, stats = run_stage(reads, k, dmin, **kw)
    check_against_golden(st.result(False), st.result(True), d, f"{os.path.basename(path)} {variant}")
    st.close()


def test_lookup_and_dump(tmp_path):
    d, k, dmin, reads, _ = load_golden([p for p in GOLDEN if p.endswith("synth_k55.npz")][0])
    st, _ = run_stage(reads, k, dmin)
    c, l, r = st.lookup(d["d1_keys"][:200])
    assert np.array_equal(c, d["d1_counts"][:200])
    assert np.array_equal(l.view(np.uint8), d["d1_left"][:200]) and np.array_equal(r.view(np.uint8), d["d1_right"][:200])
    absent = d["d1_keys"][:5] ^ np.uint64(1 << 40)
    c, _, _ = st.lookup(absent)
    assert not c.any()
    path = tmp_path / "kmers.txt"
    st.dump_sorted(str(path))
    lines = path.read_text().splitlines()
    assert len(lines) == len(d["d1_counts"])
    kmer, cnt, left, right = lines[0].split()  # `KMER count L R`, kmer_dht.cpp:286
    assert len(kmer) == k and int(cnt) == int(d["d1_counts"][0]) and left == chr(d["d1_left"][0]) and right == chr(d["d1_right"][0])
    assert np.array_equal(K.kcount.pack_kmer(kmer), d["d1_keys"][0])
    st.close()


@pytest.mark.parametrize("name", ["synth_k21.npz", "synth_k33.npz", "synth_k77.npz", "synth_k99.npz"])
def test_lookup_index_finds_every_survivor_and_nothing_else(name):
    # KmerDHT::get_local_kmer_counts (kmer_dht.cpp:199-203): the entry of a surviving k-mer, nullptr otherwise
    import torch

    match = [p for p in GOLDEN if p.endswith(name)]
    if not match:
        pytest.skip(f"no golden file {name}")
    d, k, dmin, reads, _ = load_golden(match[0])
    st, _ = run_stage(reads, k, dmin)
    keys = np.ascontiguousarray(d["d1_keys"]).reshape(len(d["d1_counts"]), -1)
    n = len(keys)
    # purged / dropped keys of the extended dump (count >= 2 but X / F) and perturbed keys must miss
    rng = np.random.default_rng(7)
    fake = keys.copy()
    fake[:, -1] ^= np.uint64(1) << np.uint64(63 - 2 * ((k - 1) % 32))  # flip the last base
    have = {kk.tobytes() for kk in keys}
    fake = np.array([f for f in fake if f.tobytes() not in have], dtype=np.uint64).reshape(-1, keys.shape[1])
    batch = np.concatenate([keys, fake])
    perm = rng.permutation(len(batch))
    c, l, r = st.lookup(batch[perm])
    want_c = np.concatenate([d["d1_counts"], np.zeros(len(fake), np.uint16)])[perm]
    want_l = np.concatenate([d["d1_left"], np.zeros(len(fake), np.uint8)])[perm]
    want_r = np.concatenate([d["d1_right"], np.zeros(len(fake), np.uint8)])[perm]
    assert np.array_equal(c, want_c)
    assert np.array_equal(l.view(np.uint8), want_l) and np.array_equal(r.view(np.uint8), want_r)
    # the same batch resident in HBM
    t = torch.from_numpy(batch[perm].view(np.int64)).cuda()
    tc, tl, tr = st.lookup_device(t)
    assert np.array_equal(tc.cpu().numpy().view(np.uint16), want_c)
    assert np.array_equal(tl.cpu().numpy(), want_l) and np.array_equal(tr.cpu().numpy(), want_r)
    # sorting moves the entries: the index is rebuilt
    st.result(False)
    c2, _, _ = st.lookup(batch[perm][: min(1000, n)])
    assert np.array_equal(c2, want_c[: min(1000, n)])
    st.close()


def test_reset_and_rerun_is_idempotent():
    d, k, dmin, reads, _ = load_golden([p for p in GOLDEN if p.endswith("synth_k33.npz")][0])
    st, _ = run_stage(reads, k, dmin)
    first = st.checksum(True)
    for _ in range(2):
        st.reset()
        st.run()
        assert st.checksum(True) == first
    st.close()
