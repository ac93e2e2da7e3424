"""Host-side I/O next to the hot path (partly): FASTQ -> PackedRead bytes (kc_fastq_read, after
FastqReader::get_next_fq_record, src/fastq.cpp:1028-1120, and the PackedRead encoding, src/packed_reads.cpp:99-126) on the
CPU; the reference's per-rank gzip dump (KmerDHT::dump_kmers, src/kcount/kmer_dht.cpp:274-297) on the GPU."""
import gzip

import numpy as np
import pytest

import mhm2_kmer_analysis_v2_b200 as K
from mhm2_kmer_analysis_v2_b200 import _lib


def _reads(n=400, seed=11):
    return K.generate(None, n, seed=seed, n_genomes=2, genome_len_min=3000, genome_len_max=6000, read_len_min=30, n_rate=0.01)


@pytest.mark.parametrize("ext", [".fq", ".fq.gz"])
def test_fastq_round_trip(tmp_path, ext):
    r = _reads()
    p = tmp_path / ("reads" + ext)
    r.to_fastq(p)
    q = K.ReadSet.from_fastq(p)
    assert len(q) == len(r)
    assert np.array_equal(q.bytes, r.bytes) and np.array_equal(q.offsets, r.offsets) and np.array_equal(q.lens, r.lens)
    few = K.ReadSet.from_fastq(p, max_reads=7)
    assert len(few) == 7 and np.array_equal(few.bytes, r.slice(0, 7).bytes)


def test_fastq_encoding_details(tmp_path):
    p = tmp_path / "x.fq"
    # lower case, IUPAC codes and N -> code 4 / bases; quality clamps at 31 above the offset and at 0 below it; CRLF; blank line
    p.write_bytes(b"@a\nacgtNRYn\n+\n!!5I~II#\r\n\n@b desc\nGG\n+b\nI!\n")
    r = K.ReadSet.from_fastq(p, qual_offset=33)
    assert list(r.lens) == [8, 2] and list(r.offsets) == [0, 8]
    assert [int(x) & 7 for x in r.bytes] == [0, 1, 2, 3, 4, 4, 4, 4, 2, 2]
    assert [int(x) >> 3 for x in r.bytes] == [0, 0, 20, 31, 31, 31, 31, 2, 31, 0]
    r64 = K.ReadSet.from_fastq(p, qual_offset=64)
    assert [int(x) >> 3 for x in r64.bytes][:4] == [0, 0, 0, 9]


@pytest.mark.parametrize("body", [b"a\nACGT\n+\nIIII\n", b"@a\nACGT\n-\nIIII\n", b"@a\nACGT\n+\nIII\n", b"@a\nACGT\n+\n"])
def test_fastq_rejects_malformed_records(tmp_path, body):
    p = tmp_path / "bad.fq"
    p.write_bytes(body)
    with pytest.raises(_lib.KcError):
        K.ReadSet.from_fastq(p)
    with pytest.raises(_lib.KcError):
        K.ReadSet.from_fastq(tmp_path / "missing.fq")


@pytest.mark.gpu
def test_gz_dump_like_the_reference(oracle_mod, tmp_path):
    """kc_dump_kmers: unordered `KMER count L R` lines, gzipped -- same multiset of lines as the oracle's survivors; and the
    stage fed from a FASTQ file gives the table of the stage fed from memory."""
    k = 33
    r = _reads(4000, seed=21)
    fq = tmp_path / "reads.fq.gz"
    r.to_fastq(fq)
    st = K.KcountStage(k)
    st.add_reads(K.ReadSet.from_fastq(fq))
    st.run()
    o = oracle_mod.Oracle(k, slots_per_rank=4 * r.counted_kmers(k))
    o.add_reads(r.bytes, r.offsets, r.lens)
    o.finalize()
    assert st.checksum(False) == o.checksum(False)
    out = tmp_path / f"kmers-{k}.txt.gz"
    st.dump_kmers(str(out))
    with gzip.open(out, "rt") as f:
        lines = sorted(f.read().splitlines())
    ref = o.result(False)
    want = sorted(f"{s} {c} {l.decode()} {rr.decode()}" for s, c, l, rr in zip(K.KmerResult(ref.keys, ref.counts, ref.left, ref.right).kmer_strings(k),
                                                                        ref.counts, ref.left, ref.right))
    assert lines == want
    st.close()
