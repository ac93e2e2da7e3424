"""ctypes binding of the C ABI in include/kcount_b200.h (libkcount_b200.so, built in-tree for sm_100a).

There is no fallback: if the shared library is missing the import fails with instructions to build it, and every
compute entry point fails with KC_ERR_CUDA when no CUDA device is present.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkcount_b200.so")

KC_OK = 0
KC_ERR_ARG, KC_ERR_CUDA, KC_ERR_NOMEM, KC_ERR_TABLE_FULL, KC_ERR_BINS_FULL, KC_ERR_STATE, KC_ERR_IO = -1, -2, -3, -4, -5, -6, -7
KC_ABI_VERSION = 2
KC_XCHG_PULL, KC_XCHG_COPY, KC_MAX_SLICES = 0, 1, 12
KC_FLAG_KEEP_EXT_COUNTS = 1
KC_FLAG_NO_WARP_DEDUP = 2
KC_FLAG_GLOBAL_TABLE = 4

u8p, u16p, u64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint16), C.POINTER(C.c_uint64)


class KcConfig(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32), ("k", C.c_uint32), ("m", C.c_uint32), ("qual_offset", C.c_int32),
        ("qual_cutoff", C.c_int32), ("dmin_thres", C.c_int32), ("dyn_min_depth", C.c_double), ("use_filter", C.c_int32),
        ("rank", C.c_int32), ("n_ranks", C.c_int32), ("device", C.c_int32), ("est_kmers", C.c_uint64),
        ("sequencing_depth", C.c_int32), ("base_error_rate", C.c_double), ("table_slots", C.c_uint64),
        ("mem_budget_bytes", C.c_uint64), ("local_bins", C.c_uint32), ("record_bytes", C.c_uint32), ("flags", C.c_uint32),
        ("reserved", C.c_uint32 * 7),
    ]


class KcStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "n_reads", "n_bases", "raw_kmers", "counted_kmers", "supermers_out", "supermer_bytes_out", "supermers_in",
        "inserted_kmers", "unique_kmers", "filtered_singletons", "ge2_kmers", "final_kmers", "purged_kmers", "sum_counts",
        "max_count", "dropped", "table_slots", "table_bytes", "filter_bytes")] + [
        ("avg_kmer_count", C.c_double), ("ms_parse", C.c_double), ("ms_insert", C.c_double), ("ms_finalize", C.c_double),
        ("kernel_launches", C.c_uint64), ("ms_parse_kernel", C.c_double), ("ms_insert_kernel", C.c_double),
        ("ms_table_clear", C.c_double), ("insert_kernel_launches", C.c_uint64), ("local_bins", C.c_uint64),
        ("split_passes", C.c_uint64), ("result_bytes", C.c_uint64), ("ms_refine_kernel", C.c_double), ("ms_exchange_copy", C.c_double),
        ("exchange_copy_bytes", C.c_uint64), ("pipeline_slices", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class KcSynthConfig(C.Structure):
    _fields_ = [
        ("seed", C.c_uint64), ("n_genomes", C.c_uint32), ("genome_len_min", C.c_uint64), ("genome_len_max", C.c_uint64),
        ("abundance_sigma", C.c_double), ("read_len", C.c_uint32), ("read_len_min", C.c_uint32), ("sub_rate", C.c_double),
        ("lowq_rate", C.c_double), ("lowq_rate_sub", C.c_double), ("n_rate", C.c_double), ("paired", C.c_int32),
        ("insert_mean", C.c_double), ("insert_sd", C.c_double), ("qual_offset", C.c_int32), ("reserved", C.c_uint32 * 5),
    ]


# every symbol include/kcount_b200.h declares: (restype, argtypes)
vp = C.c_void_p
PROTOTYPES = {
    "kc_config_default": (None, [C.POINTER(KcConfig)]),
    "kc_create": (C.c_int, [C.POINTER(KcConfig), C.POINTER(vp)]),
    "kc_destroy": (None, [vp]),
    "kc_last_error": (C.c_char_p, [vp]),
    "kc_version": (C.c_char_p, []),
    "kc_device_count": (C.c_int, []),
    "kc_add_reads": (C.c_int, [vp, vp, C.c_uint64, vp, vp, C.c_uint64]),
    "kc_add_packed_reads": (C.c_int, [vp, vp, C.c_uint64, C.c_uint32]),
    "kc_clear_reads": (C.c_int, [vp]),
    "kc_parse": (C.c_int, [vp]),
    "kc_exchange_send_sizes": (C.c_int, [vp, vp]),
    "kc_exchange_send_ptr": (C.c_int, [vp, C.c_int, C.POINTER(vp)]),
    "kc_exchange_recv_buffer": (C.c_int, [vp, vp, C.POINTER(vp)]),
    "kc_exchange_recv_offsets": (C.c_int, [vp, vp]),
    "kc_exchange_done": (C.c_int, [vp]),
    "kc_exchange_export": (C.c_int, [vp, vp, vp, vp]),
    "kc_exchange_import": (C.c_int, [vp, C.c_int, vp, C.c_uint64, C.c_uint64]),
    "kc_exchange_release": (C.c_int, [vp]),
    "kc_exchange_set_mode": (C.c_int, [vp, C.c_int]),
    "kc_exchange_export_slices": (C.c_int, [vp, C.POINTER(C.c_uint32), vp]),
    "kc_exchange_import_slices": (C.c_int, [vp, C.c_int, C.c_uint32, vp]),
    "kc_insert": (C.c_int, [vp]),
    "kc_finalize": (C.c_int, [vp, C.POINTER(KcStats)]),
    "kc_run": (C.c_int, [vp, C.POINTER(KcStats)]),
    "kc_reset": (C.c_int, [vp]),
    "kc_sync": (C.c_int, [vp]),
    "kc_get_stats": (C.c_int, [vp, C.POINTER(KcStats)]),
    "kc_timer_mark": (C.c_int, [vp, C.c_int]),
    "kc_timer_elapsed_ms": (C.c_int, [vp, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "kc_key_words": (C.c_uint32, [vp]),
    "kc_result_size": (C.c_int, [vp, C.c_int, C.POINTER(C.c_uint64)]),
    "kc_result_sort": (C.c_int, [vp, C.c_int]),
    "kc_result_copy": (C.c_int, [vp, C.c_int, vp, vp, vp, vp, vp]),
    "kc_set_result_sink": (C.c_int, [vp, vp, vp, vp, vp, C.c_uint64]),
    "kc_result_device_ptrs": (C.c_int, [vp, C.c_int, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
    "kc_result_checksum": (C.c_int, [vp, C.c_int, C.POINTER(C.c_uint64)]),
    "kc_dump_sorted": (C.c_int, [vp, C.c_char_p, C.c_int]),
    "kc_dump_kmers": (C.c_int, [vp, C.c_char_p, C.c_int]),
    "kc_lookup": (C.c_int, [vp, vp, C.c_uint64, vp, vp, vp]),
    "kc_checksum_noalloc": (C.c_int, [C.c_int, vp, vp, vp, vp, vp, C.c_uint64, C.c_uint32, C.POINTER(C.c_uint64)]),
    "kc_merge_pairs": (C.c_int, [C.c_int, vp, vp, vp, vp, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.POINTER(C.c_uint64), vp]),
    "kc_add_read_pairs": (C.c_int, [vp, vp, vp, vp, vp, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, vp]),
    "kc_owner_of": (C.c_int, [vp, vp]),
    "kc_minimizer_len": (C.c_uint32, [vp]),
    "kc_host_alloc": (C.c_int, [C.POINTER(vp), C.c_size_t]),
    "kc_host_free": (C.c_int, [vp]),
}

class KcMergeStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("pairs", "merged", "ambiguous", "merged_len", "overlap_len", "max_read_len", "dropped_short",
                                          "too_long", "sum_read_lens", "out_bytes", "launches")] + [("ms_kernel", C.c_double)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class KcPairSet(C.Structure):
    _fields_ = [("seq", C.POINTER(C.c_char)), ("qual", C.POINTER(C.c_char)), ("n_bytes", C.c_uint64), ("offsets", C.POINTER(C.c_uint64)),
                ("lens", C.POINTER(C.c_uint16)), ("n_reads", C.c_uint64), ("missing_read1", C.c_uint64), ("missing_read2", C.c_uint64)]


class KcReadSet(C.Structure):
    _fields_ = [("bytes", C.POINTER(C.c_uint8)), ("n_bytes", C.c_uint64), ("offsets", C.POINTER(C.c_uint64)), ("lens", C.POINTER(C.c_uint16)),
                ("n_reads", C.c_uint64)]


# every symbol include/kc_synth.h declares (libkc_synth.so: host-only C, no CUDA -- the CPU arms of bench.py load only this)
SYNTH_PROTOTYPES = {
    "kc_synth_config_default": (None, [C.POINTER(KcSynthConfig)]),
    "kc_synth_create": (C.c_int, [C.POINTER(KcSynthConfig), C.POINTER(vp)]),
    "kc_synth_destroy": (None, [vp]),
    "kc_synth_generate": (C.c_int, [vp, C.c_uint64, C.c_uint64, vp, vp, vp, C.POINTER(C.c_uint64), C.c_int]),
    "kc_reads_save": (C.c_int, [C.c_char_p, vp, C.c_uint64, vp, vp, C.c_uint64]),
    "kc_fastq_read": (C.c_int, [C.c_char_p, C.c_int, C.c_uint64, C.POINTER(KcReadSet)]),
    "kc_read_set_free": (None, [C.POINTER(KcReadSet)]),
    "kc_fastq_read_pairs": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_uint64, C.POINTER(KcPairSet)]),
    "kc_pair_set_free": (None, [C.POINTER(KcPairSet)]),
    "kc_reads_to_packed_layout": (C.c_int, [vp, vp, vp, C.c_uint64, vp, C.c_uint64, C.c_uint64, vp, C.POINTER(C.c_uint64)]),
}
SYNTH_LIB_PATH = os.path.join(HERE, "libkc_synth.so")

_lib = None
_synth = None


def synth_lib():
    """libkc_synth.so: the seeded read generator and the read-file writer (include/kc_synth.h)."""
    global _synth
    if _synth is None:
        if not os.path.exists(SYNTH_LIB_PATH):
            raise ImportError(f"{SYNTH_LIB_PATH} is missing: build it with `make -C {os.path.join(HERE, 'csrc')}`")
        L = C.CDLL(SYNTH_LIB_PATH)
        for name, (res, args) in SYNTH_PROTOTYPES.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _synth = L
    return _synth


def build():
    """Compile the library in-tree (nvcc, sm_100a). Used by __graft_entry__.build()."""
    import subprocess

    subprocess.run(["make", "-s", "-j8", "-C", os.path.join(HERE, "csrc")], check=True)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C {os.path.join(HERE, 'csrc')}` "
                "(python -c 'import __graft_entry__ as g; g.build()'). There is no CPU fallback."
            )
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            f = getattr(L, name)  # AttributeError = the library does not export what the header declares
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


class KcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"kcount_b200 error {code}: {msg}")
        self.code = code
