"""GPU parity: the CUDA stage (through the C ABI) against the CPU oracle on the same seeded reads. Bit-exact:
sorted survivors (dump D1) and all entries with count >= 2 incl. the 8 extension counters and X/F (dump D2)."""
import numpy as np
import pytest

import mhm2_kmer_analysis_v2_b200 as K
from util import assert_same, gpu_multi, gpu_single, merge_results, oracle_run

pytestmark = pytest.mark.gpu

SMALL = dict(n_genomes=3, genome_len_min=30_000, genome_len_max=60_000, abundance_sigma=1.0, sub_rate=0.005)


@pytest.mark.parametrize("k", [21, 33, 55, 77, 99])
def test_single_gpu_matches_oracle(oracle_mod, k):
    reads = K.generate(None, 12_000, seed=100 + k, **SMALL)
    o = oracle_run(oracle_mod, reads, k)
    st, stats = gpu_single(reads, k)
    assert stats["counted_kmers"] == o.counted_kmers() == reads.counted_kmers(k)
    assert stats["inserted_kmers"] == o.counted_kmers()
    assert stats["unique_kmers"] == o.num_entries()
    assert_same(st.result(True), o.result(True), f"k={k} D2")
    assert_same(st.result(False), o.result(False), f"k={k} D1")
    assert st.checksum(False) == o.checksum(False)
    assert st.checksum(True) == o.checksum(True)
    assert stats["final_kmers"] == len(o.result(False))
    assert abs(stats["avg_kmer_count"] - o.avg_kmer_count()) < 1e-9
    st.close()


@pytest.mark.parametrize("k,G", [(21, 2), (21, 3), (33, 4), (55, 8), (77, 2), (99, 3)])
def test_owner_partition_matches_oracle(oracle_mod, k, G):
    """G emulated ranks: per-owner tables equal the oracle's per-rank tables; the union equals the 1-rank result."""
    reads = K.generate(None, 8_000, seed=200 + k, **SMALL)
    o = oracle_run(oracle_mod, reads, k, n_ranks=G)
    stages, stats = gpu_multi(reads, k, G)
    for r, st in enumerate(stages):
        assert stats[r]["supermers_in"] >= 0
        assert_same(st.result(True), o.result(True, rank=r), f"k={k} G={G} rank {r} D2")
        assert_same(st.result(False), o.result(False, rank=r), f"k={k} G={G} rank {r} D1")
    union = merge_results([st.result(True) for st in stages], stages[0].N)
    assert_same(union, o.result(True), f"k={k} G={G} union")
    assert sum(s["inserted_kmers"] for s in stats) == reads.counted_kmers(k)
    for st in stages:
        st.close()


@pytest.mark.parametrize("k,bins,rec", [(21, 16, 16), (21, 64, 0), (33, 256, 32), (55, 8, 64), (21, 1, 16), (33, 1, 32),
                                        (33, 100_000, 32), (21, 65_536, 16), (77, 200_000, 64)])
def test_local_bins_and_record_sizes(oracle_mod, k, bins, rec):
    reads = K.generate(None, 8_000, seed=300 + k, **SMALL)
    o = oracle_run(oracle_mod, reads, k)
    st, stats = gpu_single(reads, k, local_bins=bins, record_bytes=rec)
    assert_same(st.result(True), o.result(True), f"k={k} bins={bins} rec={rec}")
    st.close()


@pytest.mark.parametrize("k,G", [(33, 3), (55, 2)])
def test_two_level_partition_with_owners(oracle_mod, k, G):
    """>= 65536 fine bins: the parse routes by coarse bin, kc_refine.cu merges the G sources and sorts by fine bin."""
    reads = K.generate(None, 8_000, seed=500 + k, **SMALL)
    o = oracle_run(oracle_mod, reads, k, n_ranks=G)
    stages, stats = gpu_multi(reads, k, G, local_bins=80_000)
    for r, st in enumerate(stages):
        assert stats[r]["local_bins"] == 80_000
        assert_same(st.result(True), o.result(True, rank=r), f"k={k} G={G} rank {r} D2")
    assert sum(s["inserted_kmers"] for s in stats) == reads.counted_kmers(k)
    for st in stages:
        st.close()


@pytest.mark.parametrize("k", [21, 33, 77])
def test_filter_on_equals_filter_off(oracle_mod, k):
    reads = K.generate(None, 8_000, seed=400 + k, **SMALL)
    o = oracle_run(oracle_mod, reads, k)
    st, stats = gpu_single(reads, k, use_filter=True)
    assert_same(st.result(True), o.result(True), f"k={k} filter on")
    assert stats["unique_kmers"] < o.num_entries(), "the filter should keep most singletons out of the table"
    st.close()


def test_no_dedup_path(oracle_mod):
    reads = K.generate(None, 6_000, seed=77, **SMALL)
    o = oracle_run(oracle_mod, reads, 21)
    st, _ = gpu_single(reads, 21, warp_dedup=False)
    assert_same(st.result(True), o.result(True), "no dedup")
    st.close()


@pytest.mark.parametrize("k,bins", [(33, 100_000), (21, 0), (55, 4096)])
def test_result_sink_streams_the_survivors(oracle_mod, k, bins):
    """kc_set_result_sink: the survivors land in host arrays during kc_insert / kc_finalize (sliced count launches)."""
    reads = K.generate(None, 12_000, seed=600 + k, **SMALL)
    o = oracle_run(oracle_mod, reads, k)
    ref = o.result(False)
    st = K.KcountStage(k, local_bins=bins)
    N = st.N
    cap = len(ref) + 100
    keys, counts = np.zeros(cap * N, np.uint64), np.zeros(cap, np.uint16)
    left, right = np.zeros(cap, np.uint8), np.zeros(cap, np.uint8)
    st.set_result_sink(keys, counts, left, right)
    st.add_reads(reads)
    stats = st.run()
    n = stats["final_kmers"]
    assert n == len(ref)
    kk = keys[: n * N].reshape(n, N)
    order = np.lexsort([kk[:, i] for i in reversed(range(N))])
    assert np.array_equal(kk[order], ref.keys.reshape(n, N))
    assert np.array_equal(counts[:n][order], ref.counts)
    assert np.array_equal(left[:n][order], ref.left.view(np.uint8)) and np.array_equal(right[:n][order], ref.right.view(np.uint8))
    # a sink that is too small is an error, not a truncation
    st2 = K.KcountStage(k, local_bins=bins)
    st2.set_result_sink(keys[: 10 * N], counts[:10], left[:10], right[:10])
    st2.add_reads(reads)
    with pytest.raises(K.KcError):
        st2.run()
    st.close()
    st2.close()


@pytest.mark.parametrize("order", ["ascending", "shuffled"])
def test_chunked_input_copies(oracle_mod, monkeypatch, order):
    """A host source is copied in sub-chunks (own events, parse starts on the first one, bin capacities sampled from it);
    reads that do not lie in the byte buffer in order fall back to one chunk."""
    monkeypatch.setenv("KC_INPUT_CHUNK_BYTES", "65536")
    k = 33
    reads = K.generate(None, 12_000, seed=777, **SMALL)
    if order == "shuffled":
        perm = np.random.default_rng(5).permutation(len(reads))
        reads = K.ReadSet(reads.bytes, reads.offsets[perm].copy(), reads.lens[perm].copy())
    o = oracle_run(oracle_mod, reads, k)
    st, stats = gpu_single(reads, k)
    assert stats["inserted_kmers"] == reads.counted_kmers(k)
    assert_same(st.result(True), o.result(True), f"chunked input, {order}")
    st.close()


@pytest.mark.parametrize("k,read_len", [(21, 400), (33, 251), (77, 600)])
def test_long_reads_span_several_segments(oracle_mod, k, read_len):
    """Reads longer than the parse kernel's 160-base segment are walked in several passes per lane."""
    reads = K.generate(None, 3_000, seed=900 + k, read_len=read_len, **SMALL)
    o = oracle_run(oracle_mod, reads, k)
    st, stats = gpu_single(reads, k)
    assert stats["inserted_kmers"] == reads.counted_kmers(k)
    assert_same(st.result(True), o.result(True), f"k={k} read_len={read_len}")
    st.close()


@pytest.mark.parametrize("k,block", [(21, 1 << 16), (33, 4 << 20), (55, 1 << 14)])
def test_packed_reads_container_as_input(oracle_mod, k, block):
    """kc_add_packed_reads: the reference's in-memory PackedReads (16-byte PackedRead headers pointing into bump-allocated
    blocks, src/packed_reads.hpp:68-79, 136) goes in as it is; same table as the oracle's, same as the flat input."""
    reads = K.generate(None, 30_000, seed=900 + k, n_genomes=3, genome_len_min=30_000, genome_len_max=60_000, abundance_sigma=1.0,
                       sub_rate=0.005, read_len_min=60, n_rate=0.001)
    layout = K.PackedReadsLayout(reads, block_bytes=block)
    assert layout.n_blocks > 1 or block >= (4 << 20)
    o = oracle_run(oracle_mod, reads, k)
    st = K.KcountStage(k, keep_ext_counts=True)
    st.add_packed_reads(layout)
    st.add_packed_reads(layout)  # twice, then cleared: containers append like kc_add_reads
    st.clear_reads()
    st.add_packed_reads(layout)
    stats = st.run()
    assert stats["n_reads"] == len(reads) and stats["counted_kmers"] == reads.counted_kmers(k)
    assert_same(st.result(True), o.result(True), f"packed reads k={k}")
    assert st.checksum(False) == o.checksum(False)
    st.close()


@pytest.mark.parametrize("k,sms", [(33, 16), (77, 40)])
def test_sm_partitioned_pipeline(oracle_mod, monkeypatch, k, sms):
    """The insert pipeline of >= 4 ranks: ONE refine launch on `sms` SMs of its own beside ONE count launch on
    the others, the count warps waiting on the per-slice counters the refine kernel bumps. Forced here on one GPU
    (KC_REFINE_SMS); same table as the oracle's, pass after pass."""
    monkeypatch.setenv("KC_REFINE_SMS", str(sms))
    reads = K.generate(None, 60_000, seed=1200 + k, n_genomes=4, genome_len_min=100_000, genome_len_max=200_000, abundance_sigma=1.0,
                       sub_rate=0.005)
    o = oracle_run(oracle_mod, reads, k)
    st = K.KcountStage(k, keep_ext_counts=True, local_bins=200_000)
    st.add_reads(reads)
    for _ in range(3):
        st.reset()
        stats = st.run()
        assert stats["local_bins"] >= 65536 and stats["pipeline_slices"] == 1
        assert st.checksum(True) == o.checksum(True) and st.checksum(False) == o.checksum(False)
    assert_same(st.result(True), o.result(True), f"partitioned k={k}")
    st.close()


def test_pinned_source_may_be_overwritten_after_sync(oracle_mod):
    """include/kcount_b200.h: the source buffers of kc_add_reads "may be released or overwritten after kc_sync()". The copies
    run asynchronously on the input stream when the source is pinned, so kc_sync has to wait for that stream as well."""
    from mhm2_kmer_analysis_v2_b200.reads import PinnedBuffer

    k = 33
    reads = K.generate(None, 200_000, seed=77, **SMALL)
    o = oracle_run(oracle_mod, reads, k)
    n, nb = len(reads), len(reads.bytes)
    pb, po, pl = PinnedBuffer(nb), PinnedBuffer(n * 8), PinnedBuffer(n * 2)
    pb.array[:nb] = reads.bytes
    po.view(np.uint64, n)[:] = reads.offsets
    pl.view(np.uint16, n)[:] = reads.lens
    st = K.KcountStage(k)
    st.add_reads_ptr(pb.ptr.value, nb, po.ptr.value, pl.ptr.value, n)
    st.sync()
    pb.array[:nb] = 0xFF  # the caller reuses its buffers
    po.view(np.uint64, n)[:] = 0
    pl.view(np.uint16, n)[:] = 0
    stats = st.run()
    assert stats["counted_kmers"] == reads.counted_kmers(k)
    assert st.checksum(False) == o.checksum(False)
    st.close()
