"""Read-pair merging: kc_fastq_read_pairs + the merge kernel against the REFERENCE's merge_reads().

tests/golden/merge/* were written by tests/golden/make_merge_golden.py from oracle/_ref/merge_ref = the reference's unmodified
src/merge_reads.cpp / fastq.cpp / packed_reads.cpp behind oracle/merge_shim. Bit-exact: every base and every 5-bit quality.
  * CPU (not gpu): the kernel's per-pair arithmetic (csrc/kc_merge_core.cuh, the same source the device compiles) run by
    tests/cpp/merge_core_host.cpp behind the product's pair reader -- holds the arithmetic and the reader to the goldens here;
  * GPU: kc_merge_pairs / kc_add_read_pairs through the C ABI.
"""
import gzip
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

import mhm2_kmer_analysis_v2_b200 as K
from mhm2_kmer_analysis_v2_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "merge")
INDEX = json.load(open(os.path.join(GOLD, "index.json")))
NAMES = sorted(INDEX)


def expected(name):
    return gzip.open(os.path.join(GOLD, f"{name}.expect.txt.gz"), "rb").read()


def paths(name):
    f = [os.path.join(GOLD, x) for x in INDEX[name]["files"]]
    return f[0], (f[1] if len(f) > 1 else None)


def as_text(reads, qual_offset):
    out = []
    for i in range(len(reads)):
        b = reads.bytes[int(reads.offsets[i]):int(reads.offsets[i]) + int(reads.lens[i])]
        out.append("".join("ACGTNNNN"[x & 7] for x in b))
        out.append("".join(chr(qual_offset + (x >> 3)) for x in b))
    return ("\n".join(out) + "\n").encode() if out else b""


@pytest.fixture(scope="module")
def host_tool(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("merge") / "merge_core_host")
    pkg = os.path.join(ROOT, "mhm2_kmer_analysis_v2_b200")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I" + os.path.join(ROOT, "include"), "-o", exe, os.path.join(ROOT, "tests", "cpp", "merge_core_host.cpp"),
                    "-L" + pkg, "-lkc_synth", "-Wl,-rpath," + pkg], check=True)
    return exe


def test_goldens_are_intact():
    for n in NAMES:
        assert hashlib.sha256(expected(n)).hexdigest() == INDEX[n]["sha256"], n


@pytest.mark.parametrize("name", NAMES)
def test_merge_arithmetic_on_host_equals_reference(name, host_tool, tmp_path):
    e = INDEX[name]
    p1, p2 = paths(name)
    out = tmp_path / "o.txt"
    subprocess.run([host_tool, str(e["qual_offset"]), str(e["min_kmer_len"]), "1" if e["paired"] else "0", str(out), p1] + ([p2] if p2 else []),
                   check=True, stderr=subprocess.DEVNULL)
    assert out.read_bytes() == expected(name)


@pytest.mark.parametrize("name", NAMES)
def test_merge_oracle_equals_reference(name):
    """oracle/merge_oracle.c (plain-C restatement, the checker that travels) against the reference's own output."""
    import oracle as O

    e = INDEX[name]
    p1, p2 = paths(name)
    ps = K.PairSet.from_fastq(p1, p2, paired=e["paired"], qual_offset=e["qual_offset"])
    ob, oo, ol, ctr = O.merge_pairs(ps.seq, ps.qual, ps.offsets, ps.lens, e["qual_offset"], e["min_kmer_len"], e["paired"])
    assert as_text(K.ReadSet(ob, oo, ol), e["qual_offset"]) == expected(name)
    assert ctr["merged"] == e["merged_or_dummy_mates"]


def seeded_pairs(seed, n, **kw):
    """mates from the generator of the golden fixtures (tests/golden/make_merge_golden.py) as a PairSet"""
    import importlib.util

    spec = importlib.util.spec_from_file_location("make_merge_golden", os.path.join(ROOT, "tests", "golden", "make_merge_golden.py"))
    M = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(M)
    seqs, quals = [], []
    for s1, q1, s2, q2 in M.pairs(seed, n, **kw):
        seqs += [s1, s2]
        quals += [q1, q2]
    lens = np.array([len(s) for s in seqs], np.uint16)
    offs = np.concatenate([[0], np.cumsum(lens.astype(np.uint64))[:-1]]).astype(np.uint64)
    return K.PairSet(np.frombuffer("".join(seqs).encode(), np.uint8).copy(), np.frombuffer("".join(quals).encode(), np.uint8).copy(), offs, lens)


SEEDED = [dict(seed=101, n=3000), dict(seed=102, n=2000, lens=((20, 300), (20, 300)), insert=(180, 90), n_rate=0.01, err=0.02),
          dict(seed=103, n=2000, period=11, err=0.005), dict(seed=104, n=1500, lens=((8, 60), (8, 60)), insert=(50, 20))]


@pytest.mark.parametrize("cfg", SEEDED, ids=lambda c: f"seed{c['seed']}")
def test_merge_host_arithmetic_equals_oracle_on_seeded_pairs(cfg, host_tool, tmp_path):
    """the kernel's arithmetic (host run) against the oracle on fresh seeded pairs, beyond the committed goldens"""
    import oracle as O

    cfg = dict(cfg)
    ps = seeded_pairs(cfg.pop("seed"), cfg.pop("n"), **cfg)
    fq = tmp_path / "p.fq"
    with open(fq, "w") as f:
        for i in range(len(ps.lens)):
            o, l = int(ps.offsets[i]), int(ps.lens[i])
            f.write(f"@p{i // 2}/{i % 2 + 1}\n{bytes(ps.seq[o:o + l]).decode()}\n+\n{bytes(ps.qual[o:o + l]).decode()}\n")
    out = tmp_path / "o.txt"
    r = subprocess.run([host_tool, "33", "21", "1", str(out), str(fq)], check=True, stderr=subprocess.PIPE, text=True)
    ob, oo, ol, ctr = O.merge_pairs(ps.seq, ps.qual, ps.offsets, ps.lens)
    assert out.read_bytes() == as_text(K.ReadSet(ob, oo, ol), 33)
    words = r.stderr.split()  # "pairs N merged N ambiguous N dropped N ..."
    assert int(words[words.index("merged") + 1]) == ctr["merged"] and int(words[words.index("ambiguous") + 1]) == ctr["ambiguous"]


def test_pair_reader_details(tmp_path):
    # unpaired file: every read gets the dummy mate N with quality = offset (src/merge_reads.cpp:352-371)
    p = tmp_path / "u.fq"
    p.write_text("@a\nACGT\n+\nIIII\n@b\nGG\n+\n#I\n")
    ps = K.PairSet.from_fastq(str(p), paired=False)
    assert len(ps) == 2 and list(ps.lens) == [4, 1, 2, 1]
    assert bytes(ps.seq) == b"ACGTNGGN" and bytes(ps.qual) == b"IIII!#I!"
    # a read 2 without its read 1 gets a faked N as mate 1 (src/merge_reads.cpp:410-423); names in three formats
    p = tmp_path / "m.fq"
    p.write_text("@x/1\nAAAA\n+\nIIII\n@x/2\nTTTT\n+\nIIII\n@y/2\nCCCC\n+\nIIII\n@z-R1\nGGGG\n+\nIIII\n@z-R2\nGGGA\n+\nIIII\n")
    ps = K.PairSet.from_fastq(str(p))
    assert ps.missing_read1 == 1 and len(ps) == 3
    assert bytes(ps.seq) == b"AAAATTTT" + b"NCCCC" + b"GGGGGGGA" and list(ps.lens) == [4, 4, 1, 4, 4, 4]
    # malformed record
    p = tmp_path / "bad.fq"
    p.write_text("@x/1\nAAAA\n+\nIII\n")
    with pytest.raises(_lib.KcError):
        K.PairSet.from_fastq(str(p))


def test_merge_refuses_without_a_device():
    if _lib.lib().kc_device_count() > 0:
        pytest.skip("a CUDA device is present")
    ps = K.PairSet.from_fastq(paths("basic")[0], max_pairs=4)
    with pytest.raises(_lib.KcError) as ei:
        K.merge_pairs(ps)
    assert ei.value.code == _lib.KC_ERR_CUDA  # there is no CPU path


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_merge_kernel_equals_reference(name):
    e = INDEX[name]
    p1, p2 = paths(name)
    ps = K.PairSet.from_fastq(p1, p2, paired=e["paired"], qual_offset=e["qual_offset"])
    reads, st = K.merge_pairs(ps, qual_offset=e["qual_offset"], min_kmer_len=e["min_kmer_len"], paired=e["paired"])
    assert as_text(reads, e["qual_offset"]) == expected(name)
    assert len(reads) == e["reads"] and st["launches"] == 1 and st["too_long"] == 0
    assert st["merged"] == e["merged_or_dummy_mates"]
    if st["dropped_short"] == 0 and st["pairs"]:
        assert st["sum_read_lens"] // (2 * st["pairs"]) == e["avg_read_len"]  # what merge_reads() returns


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", SEEDED + [dict(seed=105, n=20000)], ids=lambda c: f"seed{c['seed']}")
def test_merge_kernel_equals_oracle_on_seeded_pairs(cfg):
    import oracle as O

    cfg = dict(cfg)
    ps = seeded_pairs(cfg.pop("seed"), cfg.pop("n"), **cfg)
    reads, st = K.merge_pairs(ps)
    ob, oo, ol, ctr = O.merge_pairs(ps.seq, ps.qual, ps.offsets, ps.lens)
    assert np.array_equal(reads.lens, ol)
    assert as_text(reads, 33) == as_text(K.ReadSet(ob, oo, ol), 33)
    assert (st["merged"], st["ambiguous"], st["merged_len"], st["overlap_len"]) == (ctr["merged"], ctr["ambiguous"], ctr["merged_len"], ctr["overlap_len"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["basic", "tiny"])  # tiny: dropped pairs stay behind as empty reads
def test_merged_pairs_feed_the_stage_without_leaving_hbm(name):
    """kc_add_read_pairs -> kc_run equals kc_merge_pairs -> kc_add_reads -> kc_run (and so the oracle on the merged reads)."""
    import ctypes as C

    e = INDEX[name]
    ps = K.PairSet.from_fastq(paths(name)[0])
    reads, _ = K.merge_pairs(ps)
    a = K.KcountStage(k=21)
    a.add_reads(reads)
    sa = a.run()
    ca = a.checksum()
    b = K.KcountStage(k=21)
    st = _lib.KcMergeStats()
    rc = b.L.kc_add_read_pairs(b.h, ps.seq.ctypes.data, ps.qual.ctypes.data, ps.offsets.ctypes.data, ps.lens.ctypes.data, len(ps), len(ps.seq),
                               33, e["min_kmer_len"], 1, C.byref(st))
    assert rc == 0 and st.merged == e["merged_or_dummy_mates"]
    assert st.dropped_short == (2 * len(ps) - e["reads"]) // 2
    sb = b.run()
    assert b.checksum() == ca and sb["final_kmers"] == sa["final_kmers"] and sb["counted_kmers"] == sa["counted_kmers"]
    import oracle as O  # the checker (tests only)

    o = O.Oracle(k=21)
    o.add_reads(reads.bytes, reads.offsets, reads.lens)
    o.finalize()
    assert o.checksum() == ca


@pytest.mark.gpu
def test_merge_rejects_bad_characters():
    seq = np.frombuffer(b"ACGTACGTACGTACGTACGTACGTACGTAC" + b"ACGTAC!TACGTACGTACGTACGTACGTAC", np.uint8).copy()
    qual = np.frombuffer(b"I" * 60, np.uint8).copy()
    ps = K.PairSet(seq, qual, np.array([0, 30], np.uint64), np.array([30, 30], np.uint16))  # (>= min_kmer_len: not dropped)
    with pytest.raises(_lib.KcError) as ei:
        K.merge_pairs(ps)
    assert ei.value.code == _lib.KC_ERR_ARG


def _read_both_ways(path_plain, path_gz, monkeypatch, paired=True, threads=5, chunk=64, path2=None):
    """the same content through the parallel reader (plain file, cut into many small ranges) and the streaming reader (gzip)"""
    def read(p, p2):
        try:
            ps = K.PairSet.from_fastq(p, p2, paired=paired)
            return (0, bytes(ps.seq), bytes(ps.qual), list(ps.offsets), list(ps.lens), ps.missing_read1, ps.missing_read2)
        except _lib.KcError as e:
            return (e.code,)

    monkeypatch.setenv("KC_HOST_THREADS", str(threads))
    monkeypatch.setenv("KC_FASTQ_PAR_MIN_BYTES", "0")
    monkeypatch.setenv("KC_FASTQ_PAR_MIN_CHUNK", str(chunk))
    par = read(path_plain[0], path_plain[1] if path2 else None)
    monkeypatch.setenv("KC_HOST_THREADS", "1")
    monkeypatch.delenv("KC_FASTQ_PAR_MIN_BYTES")
    stream = read(path_gz[0], path_gz[1] if path2 else None)
    return par, stream


@pytest.mark.parametrize("name", NAMES)
def test_parallel_pair_reader_equals_streaming_reader_on_the_goldens(name, tmp_path, monkeypatch):
    e = INDEX[name]
    gz = paths(name)
    plain = []
    for i, g in enumerate(gz):
        if g is None:
            plain.append(None)
            continue
        p = tmp_path / f"f{i}.fq"
        p.write_bytes(gzip.open(g, "rb").read())
        plain.append(str(p))
    monkeypatch.setenv("KC_FASTQ_PAR_STRICT", "1")  # well-formed files: the parallel path must be the one that answers
    for threads, chunk in ((5, 64), (3, 1000), (8, 7)):
        par, stream = _read_both_ways(plain, gz, monkeypatch, paired=e["paired"], threads=threads, chunk=chunk, path2=gz[1])
        assert par == stream and par[0] == 0


def test_parallel_pair_reader_equals_streaming_reader_on_awkward_files(tmp_path, monkeypatch):
    """random files with missing mates, blank lines, CRLF, '@' / '+' at the start of quality lines, comments, no final newline,
    and malformed records: same reads, same counters, same error code"""
    import random

    rng = random.Random(5)
    taken = 0
    for trial in range(60):
        recs = []
        n = rng.randint(0, 40)
        for i in range(n):
            L = rng.randint(0, 30)
            s = "".join(rng.choice("ACGTN") for _ in range(L))
            q1 = "".join(rng.choice("@+I5#!~") for _ in range(L))
            q2 = "".join(rng.choice("@+I5#!~") for _ in range(L))
            style = rng.choice(["/", "/", "/", " ", "-R"])
            nm = lambda w: {"/": f"@r{i}/{w}", " ": f"@r{i} {w}:N:0:AC", "-R": f"@r{i}-R{w}"}[style] + rng.choice(["", "", " more", "\t x"])
            mates = [(nm(1), s, q1), (nm(2), s[::-1], q2)]
            drop = rng.random()
            if drop < 0.08:
                mates = mates[:1]
            elif drop < 0.16:
                mates = mates[1:]
            recs += mates
        eol = rng.choice(["\n", "\n", "\r\n"])
        text = ""
        for (h, s, q) in recs:
            plus = rng.choice(["+", "+", "+" + h[1:]])
            text += h + eol + s + rng.choice(["", "", " "]) + eol + plus + eol + q + eol
            if rng.random() < 0.1:
                text += eol
        bad = rng.random()
        if bad < 0.1 and recs:
            text = text[: rng.randint(1, len(text))]  # cut anywhere
        elif bad < 0.15:
            text += "@x/1" + eol + "ACGT" + eol + "+" + eol + "III" + eol  # lengths differ
        elif bad < 0.2 and text.endswith(eol):
            text = text[: -len(eol)]  # no final newline
        p = tmp_path / f"t{trial}.fq"
        p.write_bytes(text.encode())
        g = tmp_path / f"t{trial}.fq.gz"
        with gzip.open(g, "wb") as f:
            f.write(text.encode())
        for paired in (True, False):
            par, stream = _read_both_ways([str(p), None], [str(g), None], monkeypatch, paired=paired, threads=rng.randint(2, 9), chunk=rng.choice([1, 16, 100, 1000]))
            assert par == stream, (trial, paired, text[:200])
        monkeypatch.setenv("KC_FASTQ_PAR_STRICT", "1")
        strict, _ = _read_both_ways([str(p), None], [str(g), None], monkeypatch, threads=4, chunk=16)
        monkeypatch.delenv("KC_FASTQ_PAR_STRICT")
        taken += strict[0] == 0
    assert taken >= 30  # most of these files are well-formed enough for the parallel path (the rest fell back)
