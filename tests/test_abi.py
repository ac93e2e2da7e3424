"""CPU: the C-ABI library loads, exports every symbol include/kcount_b200.h declares, validates its arguments, and
refuses to compute without a CUDA device (there is no CPU fallback). Host-only helpers (seeded reads, read files) work."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import mhm2_kmer_analysis_v2_b200 as K
from mhm2_kmer_analysis_v2_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "kcount_b200.h")
SYNTH_HEADER = os.path.join(ROOT, "include", "kc_synth.h")


def declared_symbols(header=HEADER):
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(kc_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound():
    L = C.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/kcount_b200.h but not exported by libkcount_b200.so"
        assert n in _lib.PROTOTYPES, f"{n} has no ctypes prototype in _lib.py"
    assert set(_lib.PROTOTYPES) <= set(names), "ctypes binds symbols the header does not declare"


def test_synth_header_symbols_live_in_their_own_library():
    """include/kc_synth.h <-> libkc_synth.so; the product library exports none of them (the CPU arms of bench.py generate
    their reads without mapping libkcount_b200.so)."""
    S = C.CDLL(_lib.SYNTH_LIB_PATH)
    L = C.CDLL(_lib.LIB_PATH)
    names = declared_symbols(SYNTH_HEADER)
    assert set(names) == set(_lib.SYNTH_PROTOTYPES)
    for n in names:
        assert hasattr(S, n), f"{n} declared in include/kc_synth.h but not exported by libkc_synth.so"
        assert not hasattr(L, n), f"{n} is exported by the product library as well"


def test_struct_layouts_match_the_header():
    L = _lib.lib()
    cfg = _lib.KcConfig()
    L.kc_config_default(C.byref(cfg))
    assert (cfg.abi_version, cfg.k, cfg.m, cfg.qual_offset, cfg.qual_cutoff, cfg.dmin_thres) == (_lib.KC_ABI_VERSION, 21, 0, 33, 20, 2)
    assert cfg.dyn_min_depth == 0.9 and cfg.n_ranks == 1 and cfg.device == -1 and cfg.sequencing_depth == 4
    assert cfg.base_error_rate == 0.005 and cfg.flags == 0 and cfg.local_bins == 0
    s = _lib.KcSynthConfig()
    _lib.synth_lib().kc_synth_config_default(C.byref(s))
    assert s.seed == 1001 and s.read_len == 150 and s.sub_rate == 0.005 and s.qual_offset == 33
    assert L.kc_version().decode().startswith("kcount_b200")


@pytest.mark.parametrize("field,value", [("k", 32), ("k", 64), ("k", 1), ("k", 128), ("abi_version", 7), ("n_ranks", 0), ("rank", 3),
                                         ("record_bytes", 24)])
def test_create_rejects_bad_arguments(field, value):
    L = _lib.lib()
    cfg = _lib.KcConfig()
    L.kc_config_default(C.byref(cfg))
    setattr(cfg, field, value)
    h = C.c_void_p()
    rc = L.kc_create(C.byref(cfg), C.byref(h))
    assert rc == _lib.KC_ERR_ARG and not h.value
    assert L.kc_last_error(None)


def test_no_cpu_fallback():
    """Without a device kc_create fails with KC_ERR_CUDA; the Python layer raises instead of computing elsewhere."""
    L = _lib.lib()
    if L.kc_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(K.KcError) as e:
        K.KcountStage(21)
    assert e.value.code == _lib.KC_ERR_CUDA and "no CPU path" in str(e.value)
    for entry in (L.kc_parse, L.kc_insert, L.kc_reset, L.kc_sync):
        assert entry(None) == _lib.KC_ERR_ARG


def test_synthetic_reads_are_seeded_and_sliceable():
    a = K.generate("C1", 300)
    b = K.generate("C1", 300)
    assert np.array_equal(a.bytes, b.bytes) and np.array_equal(a.offsets, b.offsets)
    c = K.generate("C1", 100, first=200)  # read i depends only on (seed, i)
    assert np.array_equal(c.bytes, a.slice(200, 300).bytes)
    assert len(a) == 300 and a.n_bases == 300 * 150 and a.counted_kmers(21) == 300 * 128
    assert int((a.bytes & 7).max()) <= 3, "BASELINE configs contain no N"
    q = a.bytes >> 3
    assert set(np.unique(q)) <= {10, 31}, "qualities are '+' (10) or 'I' (40, stored as min(q, 31))"
    d = K.generate("C1", 300, seed=7)
    assert not np.array_equal(a.bytes, d.bytes)
    p = K.generate("C3", 10)
    assert len(p) == 10


def test_readset_encoding_and_file_round_trip(tmp_path):
    r = K.ReadSet.from_strings(["ACGTN", "", "GG"], ["II+I#", "", "5I"])
    assert r.bytes.tolist() == [0 | 31 << 3, 1 | 31 << 3, 2 | 10 << 3, 3 | 31 << 3, 4 | 2 << 3, 2 | 20 << 3, 2 | 31 << 3]
    assert r.offsets.tolist() == [0, 5, 5] and r.lens.tolist() == [5, 0, 2]
    path = tmp_path / "reads.bin"
    r.save(str(path))
    assert path.stat().st_size > len(r.bytes)
    both = K.ReadSet.concat([r, r])
    assert len(both) == 6 and both.offsets.tolist() == [0, 5, 5, 7, 12, 12]
