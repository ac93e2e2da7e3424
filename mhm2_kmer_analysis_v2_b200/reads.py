"""Seeded synthetic PackedRead streams (kc_synth_* of include/kc_synth.h, libkc_synth.so) and the configs of BASELINE.md section 4."""
import ctypes as C
import os

import numpy as np

from . import _lib


class ReadSet:
    """Reads in the reference's in-memory encoding (src/packed_reads.cpp:99-126): one byte per base =
    base(A0 C1 G2 T3 N4) | min(qual - offset, 31) << 3, plus offsets[] / lens[] (the 16-byte PackedRead headers)."""

    def __init__(self, bytes_, offsets, lens):
        self.bytes = np.ascontiguousarray(bytes_, np.uint8)
        self.offsets = np.ascontiguousarray(offsets, np.uint64)
        self.lens = np.ascontiguousarray(lens, np.uint16)

    def __len__(self):
        return len(self.lens)

    @property
    def n_bases(self):
        return int(self.lens.astype(np.uint64).sum())

    def counted_kmers(self, k):
        """sum over reads of max(0, len - k - 1): the unit of the headline metric (kcount_cpu.cpp:84-101)."""
        l = self.lens.astype(np.int64)
        return int(np.maximum(l - k - 1, 0).sum())

    def slice(self, lo, hi):
        off = self.offsets[lo:hi]
        if len(off) == 0:
            return ReadSet(np.zeros(0, np.uint8), off, self.lens[lo:hi])
        b0 = int(off[0])
        b1 = int(self.offsets[hi - 1]) + int(self.lens[hi - 1])
        return ReadSet(self.bytes[b0:b1], off - np.uint64(b0), self.lens[lo:hi])

    def save(self, path):
        L = _lib.synth_lib()
        rc = L.kc_reads_save(os.fsencode(path), self.bytes.ctypes.data, len(self.bytes), self.offsets.ctypes.data,
                             self.lens.ctypes.data, len(self.lens))
        if rc:
            raise _lib.KcError(rc, f"kc_reads_save({path})")

    @staticmethod
    def from_fastq(path, qual_offset=33, max_reads=0):
        """FASTQ (plain or .gz) -> ReadSet, as FastqReader + the PackedRead constructor do (src/fastq.cpp:1028-1120,
        src/packed_reads.cpp:99-126). Mates stay independent reads (no merging)."""
        L = _lib.synth_lib()
        rs = _lib.KcReadSet()
        rc = L.kc_fastq_read(os.fsencode(path), qual_offset, max_reads, C.byref(rs))
        if rc:
            raise _lib.KcError(rc, f"kc_fastq_read({path})")
        try:
            n, nb = rs.n_reads, rs.n_bytes
            b = np.ctypeslib.as_array(rs.bytes, shape=(nb,)).copy() if nb else np.zeros(0, np.uint8)
            o = np.ctypeslib.as_array(rs.offsets, shape=(n,)).copy() if n else np.zeros(0, np.uint64)
            ln = np.ctypeslib.as_array(rs.lens, shape=(n,)).copy() if n else np.zeros(0, np.uint16)
        finally:
            L.kc_read_set_free(C.byref(rs))
        return ReadSet(b, o, ln)

    def to_fastq(self, path, qual_offset=33):
        """Writes the reads back as FASTQ (tests, examples): names @r<i>/1, bases ACGTN, qualities offset + the stored 5 bits."""
        import gzip

        opener = gzip.open if str(path).endswith(".gz") else open
        with opener(path, "wt") as f:
            for i in range(len(self)):
                b = self.bytes[int(self.offsets[i]):int(self.offsets[i]) + int(self.lens[i])]
                seq = "".join("ACGTNNNN"[x & 7] for x in b)
                q = "".join(chr(qual_offset + (x >> 3)) for x in b)
                f.write(f"@r{i}/1\n{seq}\n+\n{q}\n")

    @staticmethod
    def from_strings(seqs, quals=None, qual_offset=33):
        """Build a ReadSet from ACGTN strings (+ optional quality strings), as PackedRead's constructor does."""
        code = {c: i for i, c in enumerate("ACGT")}
        bs, offs, lens, at = [], [], [], 0
        for i, s in enumerate(seqs):
            q = quals[i] if quals is not None else "I" * len(s)
            assert len(q) == len(s)
            for ch, qc in zip(s.upper(), q):
                bs.append(code.get(ch, 4) | (min(ord(qc) - qual_offset, 31) << 3))
            offs.append(at)
            lens.append(len(s))
            at += len(s)
        return ReadSet(np.array(bs, np.uint8), np.array(offs, np.uint64), np.array(lens, np.uint16))

    @staticmethod
    def concat(sets):
        bs = np.concatenate([s.bytes for s in sets])
        lens = np.concatenate([s.lens for s in sets])
        base, offs = 0, []
        for s in sets:
            offs.append(s.offsets + np.uint64(base))
            base += len(s.bytes)
        return ReadSet(bs, np.concatenate(offs), lens)


class PairSet:
    """Mates of read pairs as FASTQ characters (what merging needs: full quality values): reads 2 i and 2 i + 1 are pair i.
    seq / qual are uint8 arrays of characters, offsets / lens index them. The input of merge_pairs()."""

    def __init__(self, seq, qual, offsets, lens, missing_read1=0, missing_read2=0):
        self.seq, self.qual, self.offsets, self.lens = seq, qual, offsets, lens
        self.missing_read1, self.missing_read2 = missing_read1, missing_read2

    def __len__(self):
        return len(self.lens) // 2

    @staticmethod
    def from_fastq(path, path2=None, paired=True, qual_offset=33, max_pairs=0):
        """kc_fastq_read_pairs: one interleaved file, two files of mates, or (paired=False) an unpaired file whose reads get the
        reference's dummy mate "N" (src/merge_reads.cpp:340-466)."""
        L = _lib.synth_lib()
        ps = _lib.KcPairSet()
        rc = L.kc_fastq_read_pairs(os.fsencode(path), os.fsencode(path2) if path2 else None, 1 if paired else 0, qual_offset, max_pairs,
                                   C.byref(ps))
        if rc:
            raise _lib.KcError(rc, f"kc_fastq_read_pairs({path})")
        try:
            n, nb = ps.n_reads, ps.n_bytes
            as_u8 = lambda p: np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(nb,)).copy() if nb else np.zeros(0, np.uint8)
            seq, qual = as_u8(ps.seq), as_u8(ps.qual)
            o = np.ctypeslib.as_array(ps.offsets, shape=(n,)).copy() if n else np.zeros(0, np.uint64)
            ln = np.ctypeslib.as_array(ps.lens, shape=(n,)).copy() if n else np.zeros(0, np.uint16)
            return PairSet(seq, qual, o, ln, ps.missing_read1, ps.missing_read2)
        finally:
            L.kc_pair_set_free(C.byref(ps))


def merge_pairs(pairs, qual_offset=33, min_kmer_len=21, paired=True, device=0):
    """Read-pair merging on the GPU (kc_merge_pairs; replaces the pair loop of merge_reads(), src/merge_reads.cpp:466-650):
    PairSet -> (ReadSet of the merged / unmerged reads in the PackedRead encoding, statistics). No CPU path."""
    L = _lib.lib()
    n = len(pairs)
    out_b = np.zeros(max(len(pairs.seq), 1), np.uint8)
    out_o = np.zeros(max(2 * n, 1), np.uint64)
    out_l = np.zeros(max(2 * n, 1), np.uint16)
    n_out = C.c_uint64()
    st = _lib.KcMergeStats()
    rc = L.kc_merge_pairs(device, pairs.seq.ctypes.data, pairs.qual.ctypes.data, pairs.offsets.ctypes.data, pairs.lens.ctypes.data, n,
                          len(pairs.seq), qual_offset, min_kmer_len, 1 if paired else 0, out_b.ctypes.data, out_o.ctypes.data,
                          out_l.ctypes.data, C.byref(n_out), C.byref(st))
    if rc:
        raise _lib.KcError(rc, "kc_merge_pairs")
    k = n_out.value
    return ReadSet(out_b, out_o[:k].copy(), out_l[:k].copy()), st.as_dict()


class PackedReadsLayout:
    """A ReadSet laid out as the reference's PackedReads holds it in memory (src/packed_reads.hpp:68-79, 136): a contiguous
    array of 16-byte PackedRead headers (bit fields + pointer) and the read bytes bump-allocated in 4 MB blocks. The blocks
    live in `arena` (kept alive here) in descending address order. Input of KcountStage.add_packed_reads."""

    BLOCK = 4 * 1024 * 1024

    def __init__(self, reads, block_bytes=BLOCK):
        n = len(reads)
        # a read never straddles two blocks: the tail of a block may stay empty
        per_block = max(block_bytes - int(reads.lens.max(initial=0)), 1)
        n_blocks = len(reads.bytes) // per_block + 2
        self.arena = np.empty(n_blocks * block_bytes, np.uint8)
        self.headers = np.empty((n, 2), np.uint64)
        used = C.c_uint64()
        rc = _lib.synth_lib().kc_reads_to_packed_layout(reads.bytes.ctypes.data, reads.offsets.ctypes.data, reads.lens.ctypes.data, n,
                                                        self.arena.ctypes.data, len(self.arena), block_bytes, self.headers.ctypes.data, C.byref(used))
        if rc:
            raise _lib.KcError(rc, "kc_reads_to_packed_layout")
        self.n_blocks = used.value

    def __len__(self):
        return len(self.headers)


class PinnedBuffer:
    """Page-locked host memory from kc_host_alloc (cudaHostAlloc), viewed as a numpy u8 array."""

    def __init__(self, nbytes):
        self.L = _lib.lib()
        self.ptr = C.c_void_p()
        self.nbytes = max(int(nbytes), 1)
        rc = self.L.kc_host_alloc(C.byref(self.ptr), self.nbytes)
        if rc:
            raise _lib.KcError(rc, f"kc_host_alloc({self.nbytes})")
        self.array = np.ctypeslib.as_array(C.cast(self.ptr, C.POINTER(C.c_uint8)), shape=(self.nbytes,))

    def view(self, dtype, count, offset=0):
        return self.array[offset:offset + count * np.dtype(dtype).itemsize].view(dtype)

    def free(self):
        if self.ptr:
            self.array = None
            self.L.kc_host_free(self.ptr)
            self.ptr = None

    __del__ = free


def synth_config(**kw):
    cfg = _lib.KcSynthConfig()
    _lib.synth_lib().kc_synth_config_default(C.byref(cfg))
    for k, v in kw.items():
        if not hasattr(cfg, k):
            raise TypeError(f"unknown synth option {k}")
        setattr(cfg, k, v)
    return cfg


# BASELINE.md section 4 (n_reads is the config's full size; callers may take a prefix of the stream)
CONFIGS = {
    "C1": dict(synth=dict(seed=1001, n_genomes=1, genome_len_min=5_000_000, genome_len_max=5_000_000, sub_rate=0.005),
               n_reads=1_000_000, k=(21,)),
    "C2": dict(synth=dict(seed=2002, n_genomes=100, genome_len_min=1_000_000, genome_len_max=5_000_000, abundance_sigma=1.0,
                          sub_rate=0.005), n_reads=20_000_000, k=(33,)),
    "C3": dict(synth=dict(seed=3003, n_genomes=100, genome_len_min=1_000_000, genome_len_max=5_000_000, abundance_sigma=1.0,
                          sub_rate=0.005, paired=1, insert_mean=300.0, insert_sd=30.0), n_reads=200_000_000, k=(21, 33, 55)),
    "C4": dict(synth=dict(seed=2002, n_genomes=100, genome_len_min=1_000_000, genome_len_max=5_000_000, abundance_sigma=1.0,
                          sub_rate=0.005), n_reads=20_000_000, k=(77, 99)),
    "C5": dict(synth=dict(seed=5005, n_genomes=100, genome_len_min=1_000_000, genome_len_max=5_000_000, abundance_sigma=1.0,
                          sub_rate=0.01), n_reads=20_000_000, k=(21, 33)),
}


class Synth:
    """Handle on a seeded read stream; read i depends only on (seed, i)."""

    def __init__(self, cfg=None, **kw):
        self.L = _lib.synth_lib()
        self.cfg = cfg if cfg is not None else synth_config(**kw)
        self.h = C.c_void_p()
        rc = self.L.kc_synth_create(C.byref(self.cfg), C.byref(self.h))
        if rc:
            raise _lib.KcError(rc, "kc_synth_create")

    def close(self):
        if self.h:
            self.L.kc_synth_destroy(self.h)
            self.h = None

    __del__ = close

    def generate(self, first, n, n_threads=0, out_bytes=None):
        n_threads = n_threads or min(32, os.cpu_count() or 1)
        cap = n * self.cfg.read_len
        b = out_bytes if out_bytes is not None else np.empty(cap, np.uint8)
        assert len(b) >= cap
        off = np.empty(n, np.uint64)
        ln = np.empty(n, np.uint16)
        nb = C.c_uint64()
        rc = self.L.kc_synth_generate(self.h, first, n, b.ctypes.data, off.ctypes.data, ln.ctypes.data, C.byref(nb), n_threads)
        if rc:
            raise _lib.KcError(rc, "kc_synth_generate")
        return ReadSet(b[: nb.value], off, ln)


def generate(config="C1", n_reads=None, first=0, **overrides):
    """Reads [first, first + n_reads) of a named BASELINE config (or of explicit kc_synth options)."""
    spec = dict(CONFIGS[config]["synth"]) if config in CONFIGS else {}
    spec.update(overrides)
    s = Synth(**spec)
    try:
        n = n_reads if n_reads is not None else CONFIGS[config]["n_reads"]
        return s.generate(first, n)
    finally:
        s.close()
