"""TEST INFRASTRUCTURE -- ctypes view of oracle/libkcount_oracle.so (the C restatement of the reference's
CPU kcount) and readers for the dumps written by oracle/_ref/kcount_ref (the reference's own sources).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkcount_oracle.so")
REF_BIN = os.path.join(HERE, "_ref", "kcount_ref")

_lib = None


def build():
    """Compile the C restatement (and, where /root/reference exists, the reference build)."""
    subprocess.run(["make", "-s", "-C", HERE, "libkcount_oracle.so"], check=True)
    if os.path.isdir("/root/reference/src"):
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        u64, vp, i32 = C.c_uint64, C.c_void_p, C.c_int
        L.kco_create.restype = vp
        L.kco_create.argtypes = [i32, i32, i32, i32, i32, C.c_double, i32, u64]
        L.kco_destroy.argtypes = [vp]
        L.kco_minimizer_len.argtypes = [vp]
        L.kco_key_words.argtypes = [vp]
        L.kco_add_reads.argtypes = [vp, vp, vp, vp, u64]
        L.kco_insert_supermer.argtypes = [vp, i32, C.c_char_p, i32, i32]
        L.kco_finalize.argtypes = [vp]
        for f in ("kco_num_supermers", "kco_num_entries"):
            getattr(L, f).restype = u64
            getattr(L, f).argtypes = [vp, i32]
        for f in ("kco_num_dropped", "kco_counted_kmers"):
            getattr(L, f).restype = u64
            getattr(L, f).argtypes = [vp]
        L.kco_result_size.restype = u64
        L.kco_result_size.argtypes = [vp, i32, i32]
        L.kco_result_copy.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp]
        L.kmo_merge_pairs.restype = i32
        L.kmo_merge_pairs.argtypes = [vp, vp, vp, vp, u64, i32, i32, i32, vp, vp, vp, vp, vp]
        L.kco_result_checksum.restype = u64
        L.kco_result_checksum.argtypes = [vp, i32, i32]
        L.kco_avg_kmer_count.restype = C.c_double
        L.kco_avg_kmer_count.argtypes = [vp]
        L.kco_read_supermers.argtypes = [vp, vp, i32, vp, vp, vp, i32]
        L.kco_quick_hash.restype = u64
        L.kco_quick_hash.argtypes = [u64]
        L.kco_murmur64.restype = u64
        L.kco_murmur64.argtypes = [vp, C.c_uint32]
        L.kco_pack_kmer.argtypes = [C.c_char_p, i32, vp]
        L.kco_revcomp.argtypes = [vp, i32, vp]
        L.kco_minimizer.restype = u64
        L.kco_minimizer.argtypes = [vp, i32, i32]
        L.kco_owner.argtypes = [vp, i32, i32, i32]
        L.kco_get_ext.restype = C.c_char
        L.kco_get_ext.argtypes = [vp, C.c_uint16, i32, C.c_double]
        L.kco_entry_mix.restype = u64
        L.kco_entry_mix.argtypes = [vp, i32, C.c_uint16, C.c_char, C.c_char, vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Result:
    """Sorted result arrays: keys [n, N] u64, counts [n] u16, left/right [n] S1, ext8 [n, 8] u16 or None."""

    def __init__(self, keys, counts, left, right, ext8=None):
        self.keys, self.counts, self.left, self.right, self.ext8 = keys, counts, left, right, ext8

    def __len__(self):
        return len(self.counts)

    def same_as(self, other):
        ok = (
            self.keys.shape == other.keys.shape
            and np.array_equal(self.keys, other.keys)
            and np.array_equal(self.counts, other.counts)
            and np.array_equal(self.left, other.left)
            and np.array_equal(self.right, other.right)
        )
        if ok and self.ext8 is not None and other.ext8 is not None:
            ok = np.array_equal(self.ext8, other.ext8)
        return bool(ok)

    def first_difference(self, other):
        n = min(len(self), len(other))
        for name in ("keys", "counts", "left", "right", "ext8"):
            a, b = getattr(self, name), getattr(other, name)
            if a is None or b is None:
                continue
            d = np.nonzero((a[:n] != b[:n]).reshape(n, -1).any(axis=1))[0]
            if len(d):
                i = int(d[0])
                return f"{name} differs first at entry {i}: {a[i]} vs {b[i]} (sizes {len(self)} / {len(other)})"
        return f"sizes {len(self)} / {len(other)}"


class Oracle:
    """kcount of the reference's CPU build, restated (oracle/kcount_oracle.c)."""

    def __init__(self, k, m=0, qual_offset=33, qual_cutoff=20, dmin_thres=2, dyn_min_depth=0.9, n_ranks=1, slots_per_rank=1 << 20):
        self.L = lib()
        self.h = self.L.kco_create(k, m, qual_offset, qual_cutoff, dmin_thres, dyn_min_depth, n_ranks, slots_per_rank)
        if not self.h:
            raise ValueError("kco_create failed")
        self.k, self.n_ranks = k, n_ranks
        self.N = self.L.kco_key_words(self.h)
        self.m = self.L.kco_minimizer_len(self.h)

    def close(self):
        if self.h:
            self.L.kco_destroy(self.h)
            self.h = None

    __del__ = close

    def add_reads(self, bytes_, offsets, lens):
        bytes_ = np.ascontiguousarray(bytes_, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        lens = np.ascontiguousarray(lens, np.uint16)
        rc = self.L.kco_add_reads(self.h, _p(bytes_), _p(offsets), _p(lens), len(lens))
        assert rc == 0

    def insert_supermer(self, rank, seq, count=1):
        rc = self.L.kco_insert_supermer(self.h, rank, seq.encode(), len(seq), count)
        assert rc == 0, rc

    def finalize(self):
        self.L.kco_finalize(self.h)
        if self.L.kco_num_dropped(self.h):
            raise RuntimeError("oracle table overflow: parity undefined (kcount_cpu.cpp:266)")

    def result(self, extended=False, rank=-1):
        n = self.L.kco_result_size(self.h, int(extended), rank)
        keys = np.zeros((n, self.N), np.uint64)
        counts = np.zeros(n, np.uint16)
        left = np.zeros(n, "S1")
        right = np.zeros(n, "S1")
        ext8 = np.zeros((n, 8), np.uint16) if extended else None
        self.L.kco_result_copy(self.h, int(extended), rank, _p(keys), _p(counts), _p(left), _p(right), _p(ext8))
        return Result(keys, counts, left, right, ext8)

    def checksum(self, extended=False, rank=-1):
        return self.L.kco_result_checksum(self.h, int(extended), rank)

    def num_supermers(self, rank):
        return self.L.kco_num_supermers(self.h, rank)

    def num_entries(self, rank=-1):
        return self.L.kco_num_entries(self.h, rank)

    def counted_kmers(self):
        return self.L.kco_counted_kmers(self.h)

    def avg_kmer_count(self):
        return self.L.kco_avg_kmer_count(self.h)

    def read_supermers(self, read_bytes):
        read_bytes = np.ascontiguousarray(read_bytes, np.uint8)
        cap = max(len(read_bytes), 1)
        t, f, n = (np.zeros(cap, np.int32) for _ in range(3))
        c = self.L.kco_read_supermers(self.h, _p(read_bytes), len(read_bytes), _p(t), _p(f), _p(n), cap)
        return [(int(t[i]), int(f[i]), int(n[i])) for i in range(c)]


def load_dump(path):
    """Read a D1/D2 dump written by oracle/_ref/kcount_ref (ref_harness.cpp write_part)."""
    with open(path, "rb") as f:
        hdr = np.frombuffer(f.read(32), np.uint64)
        assert hdr[0] == 0x31504D44434B, "bad dump magic"
        N, n, per = int(hdr[1]), int(hdr[2]), int(hdr[3])
        keys = np.frombuffer(f.read(8 * N * n), np.uint64).reshape(n, N).copy()
        vals = np.frombuffer(f.read(2 * per * n), np.uint16).reshape(n, per)
    if per == 3:
        return Result(keys, vals[:, 0].copy(), vals[:, 1].astype(np.uint8).view("S1"), vals[:, 2].astype(np.uint8).view("S1"))
    return Result(
        keys, vals[:, 0].copy(), vals[:, 9].astype(np.uint8).view("S1"), vals[:, 10].astype(np.uint8).view("S1"), vals[:, 1:9].copy()
    )


def result_checksum(res, extended):
    """kco_entry_mix summed over a Result (numpy restatement; used to checksum reference dumps)."""
    M1, M2 = np.uint64(0xFF51AFD7ED558CCD), np.uint64(0xC4CEB9FE1A85EC53)
    with np.errstate(over="ignore"):
        h = np.full(len(res), 0x9E3779B97F4A7C15, np.uint64)
        for i in range(res.keys.shape[1]):
            h = (h ^ res.keys[:, i]) * M1
            h ^= h >> np.uint64(32)
        v = res.counts.astype(np.uint64) | (res.left.view(np.uint8).astype(np.uint64) << np.uint64(16)) | (
            res.right.view(np.uint8).astype(np.uint64) << np.uint64(24)
        )
        h = (h ^ v) * M2
        h ^= h >> np.uint64(29)
        if extended:
            e = res.ext8.astype(np.uint64)
            e0 = e[:, 0] | (e[:, 1] << np.uint64(16)) | (e[:, 2] << np.uint64(32)) | (e[:, 3] << np.uint64(48))
            e1 = e[:, 4] | (e[:, 5] << np.uint64(16)) | (e[:, 6] << np.uint64(32)) | (e[:, 7] << np.uint64(48))
            h = (h ^ e0) * M1
            h ^= h >> np.uint64(31)
            h = (h ^ e1) * M2
            h ^= h >> np.uint64(30)
        return int(h.sum(dtype=np.uint64))


def run_reference(reads_path, k, out_prefix=None, extended=False, ranks=1, dmin=2, qual_offset=33, max_reads=0, seq_depth=4, env=None):
    """Run the reference's own CPU kcount (oracle/_ref/kcount_ref); returns its JSON summary."""
    import json

    cmd = [REF_BIN, "--reads", reads_path, "--k", str(k), "--ranks", str(ranks), "--dmin", str(dmin), "--qual-offset", str(qual_offset),
           "--seq-depth", str(seq_depth)]
    if out_prefix:
        cmd += ["--out", out_prefix]
    if extended:
        cmd += ["--extended"]
    if max_reads:
        cmd += ["--max-reads", str(max_reads)]
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run(cmd, capture_output=True, text=True, env=e)
    if r.returncode != 0:
        raise RuntimeError(f"kcount_ref failed: {r.stderr[-2000:]}")
    return json.loads(r.stdout.strip().splitlines()[-1])


def merge_pairs(seq, qual, offsets, lens, qual_offset=33, min_kmer_len=21, paired=True):
    """oracle/merge_oracle.c (the reference's merge_reads pair loop, restated): FASTQ characters of the mates (reads 2 i, 2 i + 1
    = pair i) -> (PackedRead bytes, offsets, lens of the reads the reference keeps, counters dict)."""
    seq = np.ascontiguousarray(seq, np.uint8)
    qual = np.ascontiguousarray(qual, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    lens = np.ascontiguousarray(lens, np.uint16)
    n = len(lens) // 2
    ob = np.zeros(int(lens.astype(np.uint64).sum()) + 16, np.uint8)
    oo = np.zeros(max(2 * n, 1), np.uint64)
    ol = np.zeros(max(2 * n, 1), np.uint16)
    n_out = np.zeros(1, np.uint64)
    ctr = np.zeros(4, np.uint64)
    rc = lib().kmo_merge_pairs(_p(seq), _p(qual), _p(offsets), _p(lens), n, qual_offset, min_kmer_len, int(paired), _p(ob), _p(oo), _p(ol), _p(n_out), _p(ctr))
    if rc:
        raise ValueError("merge oracle: a character the reference would DIE on")
    k = int(n_out[0])
    return ob, oo[:k].copy(), ol[:k].copy(), dict(merged=int(ctr[0]), ambiguous=int(ctr[1]), merged_len=int(ctr[2]), overlap_len=int(ctr[3]))
