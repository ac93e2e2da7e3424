"""Host-side mirror of the reference's kcount interface over the C ABI (include/kcount_b200.h).

`KcountStage` is a 1:1 object view of the ABI. `KmerDHT` / `analyze_kmers` / `PackedReads` keep the names, argument
meaning and call order of the reference (src/kcount/kmer_dht.hpp:118-182, src/kcount/kcount.hpp:70-72,
src/packed_reads.hpp:130-201) so that tests read like a caller of the reference (src/contigging.cpp:83-104).
The C++ twin of this file is include/kcount_b200.hpp.
"""
import ctypes as C
import os

import numpy as np

from . import _lib
from ._lib import KcError
from .reads import ReadSet


class KmerResult:
    """Dense result arrays: keys [n, N] u64 (reference Kmer word layout), counts [n] u16, left/right [n] S1,
    ext8 [n, 8] u16 (lA lC lG lT rA rC rG rT) when extended."""

    def __init__(self, keys, counts, left, right, ext8=None):
        self.keys, self.counts, self.left, self.right, self.ext8 = keys, counts, left, right, ext8

    def __len__(self):
        return len(self.counts)

    def kmer_strings(self, k):
        out = []
        for row in self.keys:
            out.append("".join("ACGT"[(int(row[i // 32]) >> (2 * (31 - i % 32))) & 3] for i in range(k)))
        return out


class KcountStage:
    """One k-mer analysis stage on one GPU (one rank)."""

    def __init__(self, k, m=0, qual_offset=33, qual_cutoff=20, dmin_thres=2, dyn_min_depth=0.9, use_filter=False, rank=0,
                 n_ranks=1, device=-1, est_kmers=0, sequencing_depth=4, base_error_rate=0.005, table_slots=0,
                 mem_budget_bytes=0, local_bins=0, record_bytes=0, keep_ext_counts=False, warp_dedup=True, global_table=False):
        self.L = _lib.lib()
        cfg = _lib.KcConfig()
        self.L.kc_config_default(C.byref(cfg))
        cfg.k, cfg.m, cfg.qual_offset, cfg.qual_cutoff, cfg.dmin_thres = k, m, qual_offset, qual_cutoff, dmin_thres
        cfg.dyn_min_depth, cfg.use_filter, cfg.rank, cfg.n_ranks, cfg.device = dyn_min_depth, int(use_filter), rank, n_ranks, device
        cfg.est_kmers, cfg.sequencing_depth, cfg.base_error_rate = est_kmers, sequencing_depth, base_error_rate
        cfg.table_slots, cfg.mem_budget_bytes, cfg.local_bins, cfg.record_bytes = table_slots, mem_budget_bytes, local_bins, record_bytes
        cfg.flags = ((_lib.KC_FLAG_KEEP_EXT_COUNTS if keep_ext_counts else 0) | (0 if warp_dedup else _lib.KC_FLAG_NO_WARP_DEDUP)
                     | (_lib.KC_FLAG_GLOBAL_TABLE if global_table else 0))
        self.cfg = cfg
        self.h = C.c_void_p()
        rc = self.L.kc_create(C.byref(cfg), C.byref(self.h))
        if rc:
            raise KcError(rc, self.L.kc_last_error(None).decode())
        self.k, self.n_ranks, self.rank = k, n_ranks, rank
        self.N = self.L.kc_key_words(self.h)
        self.m = self.L.kc_minimizer_len(self.h)
        self._keepalive = []

    def close(self):
        if getattr(self, "h", None):
            self.L.kc_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc:
            raise KcError(rc, self.L.kc_last_error(self.h).decode())

    # ---- input
    def add_reads(self, reads):
        """reads: ReadSet (host numpy) -- copied to the device."""
        self._ck(self.L.kc_add_reads(self.h, reads.bytes.ctypes.data, len(reads.bytes), reads.offsets.ctypes.data,
                                     reads.lens.ctypes.data, len(reads.lens)))
        self._keepalive.append(reads)  # async copies from pinned memory may still be in flight

    def add_packed_reads(self, container):
        """container: reads.PackedReadsLayout -- the reference's in-memory PackedReads (16-byte headers pointing into 4 MB
        allocator blocks). Returns when the bytes are on the device."""
        self._ck(self.L.kc_add_packed_reads(self.h, container.headers.ctypes.data, len(container), 16))

    def add_reads_ptr(self, bytes_ptr, n_bytes, offsets_ptr, lens_ptr, n_reads):
        """Raw pointers (host pinned or device), e.g. torch tensors' data_ptr()."""
        self._ck(self.L.kc_add_reads(self.h, bytes_ptr, n_bytes, offsets_ptr, lens_ptr, n_reads))

    def clear_reads(self):
        self._ck(self.L.kc_clear_reads(self.h))
        self._keepalive.clear()

    # ---- phases
    def parse(self):
        self._ck(self.L.kc_parse(self.h))

    def insert(self):
        self._ck(self.L.kc_insert(self.h))

    def finalize(self):
        st = _lib.KcStats()
        self._ck(self.L.kc_finalize(self.h, C.byref(st)))
        return st.as_dict()

    def run(self):
        st = _lib.KcStats()
        self._ck(self.L.kc_run(self.h, C.byref(st)))
        return st.as_dict()

    def reset(self):
        self._ck(self.L.kc_reset(self.h))

    def sync(self):
        self._ck(self.L.kc_sync(self.h))

    def set_result_sink(self, keys=None, counts=None, left=None, right=None):
        """Host arrays (numpy; pinned memory for full speed) that receive the survivors while the count kernel is still
        running -- the reference's device -> host hand-off in done_all_inserts. keys=None removes the sink. After
        finalize() the first stats['final_kmers'] entries are valid (unsorted)."""
        if keys is None:
            self._ck(self.L.kc_set_result_sink(self.h, None, None, None, None, 0))
            self._sink = None
            return
        cap = min(keys.size // self.N, counts.size, left.size, right.size)
        self._sink = (keys, counts, left, right)  # keep the buffers alive
        self._ck(self.L.kc_set_result_sink(self.h, keys.ctypes.data, counts.ctypes.data, left.ctypes.data, right.ctypes.data, cap))

    def stats(self):
        st = _lib.KcStats()
        self._ck(self.L.kc_get_stats(self.h, C.byref(st)))
        return st.as_dict()

    def timer_mark(self, slot):
        self._ck(self.L.kc_timer_mark(self.h, slot))

    def timer_elapsed_ms(self, a, b):
        ms = C.c_double()
        self._ck(self.L.kc_timer_elapsed_ms(self.h, a, b, C.byref(ms)))
        return ms.value

    # ---- exchange (multi-rank)
    def send_sizes(self):
        a = np.zeros(self.n_ranks, np.uint64)
        self._ck(self.L.kc_exchange_send_sizes(self.h, a.ctypes.data))
        return a

    def send_ptr(self, dst):
        p = C.c_void_p()
        self._ck(self.L.kc_exchange_send_ptr(self.h, dst, C.byref(p)))
        return p.value or 0

    def recv_buffer(self, recv_sizes):
        a = np.ascontiguousarray(recv_sizes, np.uint64)
        p = C.c_void_p()
        self._ck(self.L.kc_exchange_recv_buffer(self.h, a.ctypes.data, C.byref(p)))
        off = np.zeros(self.n_ranks, np.uint64)
        self._ck(self.L.kc_exchange_recv_offsets(self.h, off.ctypes.data))
        return (p.value or 0), off

    def export_regions(self):
        """Peer-memory exchange: (IPC handle bytes, offsets[n_ranks], used bytes[n_ranks]) of this rank's send buffer."""
        h = np.zeros(64, np.uint8)
        off, nb = np.zeros(self.n_ranks, np.uint64), np.zeros(self.n_ranks, np.uint64)
        self._ck(self.L.kc_exchange_export(self.h, h.ctypes.data, off.ctypes.data, nb.ctypes.data))
        return h, off, nb

    def export_slices(self):
        """(n_slices, offsets[n_ranks, KC_MAX_SLICES + 1]): byte offset of every pipeline-slice edge in each owner's region."""
        n = C.c_uint32()
        off = np.zeros((self.n_ranks, _lib.KC_MAX_SLICES + 1), np.uint64)
        self._ck(self.L.kc_exchange_export_slices(self.h, C.byref(n), off.ctypes.data))
        return n.value, off

    def import_slices(self, src, n_slices, offsets):
        offsets = np.ascontiguousarray(offsets, np.uint64)
        self._ck(self.L.kc_exchange_import_slices(self.h, int(src), int(n_slices), offsets.ctypes.data))

    def set_exchange_mode(self, mode):
        """'pull': the refine / count kernels read the peers' HBM themselves; 'copy': copy engines bring the peers' regions
        in slice by slice while the previous slices are refined and counted."""
        self._ck(self.L.kc_exchange_set_mode(self.h, {"pull": _lib.KC_XCHG_PULL, "copy": _lib.KC_XCHG_COPY}[mode]))

    def import_region(self, src, handle, offset, nbytes):
        handle = np.ascontiguousarray(handle, np.uint8)
        self._ck(self.L.kc_exchange_import(self.h, int(src), handle.ctypes.data, int(offset), int(nbytes)))

    def exchange_done(self):
        self._ck(self.L.kc_exchange_done(self.h))

    def exchange_release(self):
        """After the barrier that follows finalize(): frees exported send buffers that were since re-allocated."""
        self._ck(self.L.kc_exchange_release(self.h))

    # ---- results
    def result_size(self, extended=False):
        n = C.c_uint64()
        self._ck(self.L.kc_result_size(self.h, int(extended), C.byref(n)))
        return n.value

    def result(self, extended=False, sort=True):
        if sort:
            self._ck(self.L.kc_result_sort(self.h, int(extended)))
        n = self.result_size(extended)
        keys = np.zeros((n, self.N), np.uint64)
        counts = np.zeros(n, np.uint16)
        left, right = np.zeros(n, "S1"), np.zeros(n, "S1")
        ext8 = np.zeros((n, 8), np.uint16) if extended else None
        self._ck(self.L.kc_result_copy(self.h, int(extended), keys.ctypes.data, counts.ctypes.data, left.ctypes.data,
                                       right.ctypes.data, ext8.ctypes.data if extended else None))
        return KmerResult(keys, counts, left, right, ext8)

    def result_device(self, extended=False):
        """Zero-copy torch views of the result arrays in HBM (kc_result_device_ptrs): keys int64 [n, N] (the u64 words),
        counts int16 [n] (the u16 values), left / right uint8 [n]. Valid until the next count / reset / close."""
        import torch

        n = self.result_size(extended)
        ptrs = [C.c_void_p() for _ in range(5)]
        self._ck(self.L.kc_result_device_ptrs(self.h, int(extended), *[C.byref(p) for p in ptrs]))

        class _View:
            def __init__(self, ptr, shape, typestr):
                self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 2}

        dev = torch.device("cuda", torch.cuda.current_device())
        if n == 0:
            return dict(keys=torch.empty((0, self.N), dtype=torch.int64, device=dev), counts=torch.empty(0, dtype=torch.int16, device=dev),
                        left=torch.empty(0, dtype=torch.uint8, device=dev), right=torch.empty(0, dtype=torch.uint8, device=dev))
        return dict(keys=torch.as_tensor(_View(ptrs[0].value, (n, self.N), "<i8"), device=dev),
                    counts=torch.as_tensor(_View(ptrs[1].value, (n,), "<i2"), device=dev),
                    left=torch.as_tensor(_View(ptrs[2].value, (n,), "|u1"), device=dev),
                    right=torch.as_tensor(_View(ptrs[3].value, (n,), "|u1"), device=dev))

    def checksum(self, extended=False):
        v = C.c_uint64()
        self._ck(self.L.kc_result_checksum(self.h, int(extended), C.byref(v)))
        return v.value

    def dump_sorted(self, path, extended=False):
        self._ck(self.L.kc_dump_sorted(self.h, os.fsencode(path), int(extended)))

    def dump_kmers(self, path, gz=True):
        """KmerDHT::dump_kmers (src/kcount/kmer_dht.cpp:274-297): `KMER count L R` per survivor, unordered, gzip-compressed."""
        self._ck(self.L.kc_dump_kmers(self.h, os.fsencode(path), int(gz)))

    def lookup(self, keys):
        keys = np.ascontiguousarray(keys, np.uint64).reshape(-1, self.N)
        n = len(keys)
        counts, left, right = np.zeros(n, np.uint16), np.zeros(n, "S1"), np.zeros(n, "S1")
        self._ck(self.L.kc_lookup(self.h, keys.ctypes.data, n, counts.ctypes.data, left.ctypes.data, right.ctypes.data))
        return counts, left, right

    def lookup_device(self, keys):
        """kc_lookup on a batch that already lives in HBM: keys = int64 CUDA tensor [n, N] (the u64 words, bit for bit);
        returns (counts int16, left uint8, right uint8) CUDA tensors; counts are the u16 values, bit for bit."""
        import torch

        assert keys.is_cuda and keys.dtype == torch.int64 and keys.is_contiguous()
        n = keys.numel() // self.N
        counts = torch.empty(n, dtype=torch.int16, device=keys.device)
        left = torch.empty(n, dtype=torch.uint8, device=keys.device)
        right = torch.empty(n, dtype=torch.uint8, device=keys.device)
        torch.cuda.current_stream(keys.device).synchronize()  # the library queries on its own stream
        self._ck(self.L.kc_lookup(self.h, keys.data_ptr(), n, counts.data_ptr(), left.data_ptr(), right.data_ptr()))
        return counts, left, right

    def owner_of(self, key_words):
        kw = np.ascontiguousarray(key_words, np.uint64)
        return self.L.kc_owner_of(self.h, kw.ctypes.data)


# ------------------------------------------------------------------------------------------------ reference-shaped façade

_dmin_thres = 2  # process-wide in the reference too (kmer_dht.hpp:57)


def pack_kmer(s):
    """Kmer<MAX_K>::set_kmer (src/kmer.cpp:303-325): words of a k-mer string (N -> G)."""
    code = {"A": 0, "C": 1, "G": 2, "T": 3, "N": 2}
    n = (len(s) + 31) // 32
    w = [0] * n
    for i, ch in enumerate(s.upper()):
        w[i // 32] |= code[ch] << (2 * (31 - i % 32))
    return np.array(w, np.uint64)


def revcomp_kmer_string(s):
    return s.upper().translate(str.maketrans("ACGTN", "TGCAN"))[::-1]


class PackedReads:
    """The part of the reference's PackedReads a kcount caller touches (src/packed_reads.hpp:130-201)."""

    def __init__(self, qual_offset, reads=None, fname=""):
        self.qual_offset, self.fname = qual_offset, fname
        self.reads = reads if reads is not None else ReadSet(np.zeros(0, np.uint8), np.zeros(0, np.uint64), np.zeros(0, np.uint16))

    def get_local_num_reads(self):
        return len(self.reads)

    @staticmethod
    def get_total_local_num_reads(lst):
        return sum(p.get_local_num_reads() for p in lst)

    @staticmethod
    def estimate_num_kmers(kmer_len, packed_reads_list):
        """src/packed_reads.cpp:484-518: k-mers of the first 50 000 reads of each file, scaled to all reads."""
        num_kmers = num_reads = 0
        tot = PackedReads.get_total_local_num_reads(packed_reads_list)
        for pr in packed_reads_list:
            l = pr.reads.lens[:50000].astype(np.int64)
            l = l[l >= kmer_len]
            num_kmers += int((l - kmer_len + 1).sum())
            num_reads += len(l)
        return num_kmers * tot // num_reads if num_reads else 0


class KmerCounts:
    """src/kcount/kmer_dht.hpp:62-68"""
    __slots__ = ("uutig_frag", "count", "left", "right")

    def __init__(self, count, left, right):
        self.uutig_frag, self.count, self.left, self.right = None, int(count), left, right


class KmerDHT:
    """KmerDHT<MAX_K> (src/kcount/kmer_dht.hpp:118-182). One instance per rank; MAX_K is implied by k.

    The supermer aggregation store + RPC of the reference are replaced by the device-side bins and the caller's
    exchange (see mhm2_kmer_analysis_v2_b200.exchange); `local_kmers` is materialised from the dense device result only
    when somebody asks for it (the reference's eager host insert is its own bottleneck, kcount_gpu.cpp:439-471).
    """

    def __init__(self, kmer_len, my_num_kmers, my_num_ctg_kmers=0, max_kmer_store_bytes=0, max_rpcs_in_flight=100, use_qf=False,
                 sequencing_depth=4, rank=0, n_ranks=1, device=-1, **stage_options):
        if my_num_kmers <= 0:
            raise KcError(_lib.KC_ERR_ARG, "no kmers to reserve space for")  # kmer_dht.cpp:146
        self.kmer_len = kmer_len
        self.stage = KcountStage(kmer_len, use_filter=use_qf, est_kmers=int(my_num_kmers), sequencing_depth=sequencing_depth,
                                 rank=rank, n_ranks=n_ranks, device=device, **stage_options)
        self.using_ctg_kmers = False
        self._avg_kmer_count = 0.0
        self._stats = None
        self._result = None
        self._map = None

    # -- accessors, same names as the reference
    def get_minimizer_len(self):
        return self.stage.m

    def get_local_num_kmers(self):
        return self.stage.result_size(False) if self._stats else 0

    def get_num_kmers(self, all=False, group=None):
        n = self.get_local_num_kmers()
        if self.stage.n_ranks > 1:
            import torch
            import torch.distributed as dist

            t = torch.tensor([n], dtype=torch.int64, device="cuda" if dist.get_backend(group) == "nccl" else "cpu")
            dist.all_reduce(t, group=group)
            n = int(t.item())
        return n

    def get_kmer_target_rank(self, kmer, kmer_rc=None):
        """kmer: str or word array of the canonical k-mer (the reference passes the canonical one, kcount_cpu.cpp:78-87)."""
        if isinstance(kmer, str):
            rc = revcomp_kmer_string(kmer)
            w, wr = pack_kmer(kmer), pack_kmer(rc)
            kmer = wr if tuple(wr) < tuple(w) else w
        return self.stage.owner_of(kmer)

    def get_num_supermer_inserts(self):
        return self._stats["supermers_in"] if self._stats else 0

    def get_avg_kmer_count(self):
        return self._avg_kmer_count

    def _canonical_words(self, kmer):
        if isinstance(kmer, str):
            w, wr = pack_kmer(kmer), pack_kmer(revcomp_kmer_string(kmer))
            return wr if tuple(wr) < tuple(w) else w
        return np.asarray(kmer, np.uint64)

    def get_local_kmer_counts(self, kmer):
        c, l, r = self.stage.lookup(self._canonical_words(kmer))
        return KmerCounts(c[0], l[0].decode(), r[0].decode()) if c[0] else None

    def kmer_exists(self, kmer):
        return self.get_local_kmer_counts(kmer) is not None

    def get_kmer_count(self, kmer):
        kc = self.get_local_kmer_counts(kmer)
        return kc.count if kc else 0

    def add_reads(self, reads):
        self.stage.add_reads(reads)

    def flush_updates(self):
        """kmer_dht.cpp:253-258. Single rank: supermers are already binned on the device, so only the insert remains;
        multi-rank callers run exchange.exchange_supermers() between parse and this call."""
        self.stage.insert()

    def finish_updates(self):
        """kmer_dht.cpp:261-267 -> HashTableInserter::insert_into_local_hashtable."""
        self._stats = self.stage.finalize()
        self._avg_kmer_count = self._stats["avg_kmer_count"]
        self._result = None
        self._map = None

    def set_avg_kmer_count(self, v):
        self._avg_kmer_count = v

    def clear_stores(self):
        pass

    def dump_kmers(self, path=None, sorted=False):
        """kmer_dht.cpp:274-297: `KMER count L R` lines per rank, unordered, gzipped (`kmers-<k>.txt.gz`); sorted=True writes the
        plain-text sorted dump the parity tests compare instead."""
        if sorted:
            path = path or f"kmers-{self.kmer_len}.txt"
            self.stage.dump_sorted(path, False)
        else:
            path = path or f"kmers-{self.kmer_len}.txt.gz"
            self.stage.dump_kmers(path, True)
        return path

    def local_kmers(self):
        """dict kmer-string -> KmerCounts (local_kmers_begin/end)."""
        if self._map is None:
            res = self.stage.result(False)
            ks = res.kmer_strings(self.kmer_len)
            self._map = {s: KmerCounts(c, l.decode(), r.decode()) for s, c, l, r in zip(ks, res.counts, res.left, res.right)}
        return self._map

    def stats(self):
        return self._stats


def analyze_kmers(kmer_len, qual_offset, packed_reads_list, dmin_thres, kmer_dht, dump_kmers=False, group=None):
    """analyze_kmers<MAX_K> (src/kcount/kcount.cpp:143-161): count_kmers -> flush_updates -> finish_updates -> dump.

    With kmer_dht.stage.n_ranks > 1 every rank calls this collectively (torch.distributed initialised) and the supermer
    exchange replaces the reference's aggregated RPCs."""
    global _dmin_thres
    _dmin_thres = dmin_thres
    st = kmer_dht.stage
    if st.cfg.dmin_thres != dmin_thres or st.cfg.qual_offset != qual_offset or st.k != kmer_len:
        raise KcError(_lib.KC_ERR_ARG, "analyze_kmers: kmer_len / qual_offset / dmin_thres differ from the KmerDHT's stage")
    # count_kmers (kcount.cpp:55-104): hand every local read to the inserter
    for pr in packed_reads_list:
        if len(pr.reads):
            st.add_reads(pr.reads)
    st.parse()
    if st.n_ranks > 1:
        from .exchange import exchange_supermers

        exchange_supermers(st, group=group)
    kmer_dht.flush_updates()
    kmer_dht.finish_updates()
    if st.n_ranks > 1:
        from .exchange import global_avg_kmer_count

        kmer_dht.set_avg_kmer_count(global_avg_kmer_count(kmer_dht.stats(), group=group))
    if dump_kmers:
        kmer_dht.dump_kmers()
    kmer_dht.clear_stores()
    return kmer_dht.stats()
