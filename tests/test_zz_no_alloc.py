"""kc_checksum_noalloc (include/kcount_b200.h): the allocation-free variant of the result checksum -- scratch in a file-scope
device array -- must give exactly kc_result_checksum's value (integer arithmetic: tolerance 0), on two calls in a row, for the
final and the extended result, and must reject sizes beyond its stated maximum. No counterpart in the reference.
Collected last on purpose."""
import ctypes as C

import pytest

import mhm2_kmer_analysis_v2_b200 as K
from mhm2_kmer_analysis_v2_b200 import _lib


def _ptrs(stage, extended):
    p = [C.c_void_p() for _ in range(5)]
    stage._ck(stage.L.kc_result_device_ptrs(stage.h, int(extended), *[C.byref(x) for x in p]))
    return [x.value for x in p]


def _noalloc(stage, extended, n=None):
    keys, counts, left, right, ext8 = _ptrs(stage, extended)
    out = C.c_uint64(0xdeadbeef)
    rc = stage.L.kc_checksum_noalloc(0, keys, counts, left, right, ext8 if extended else None, stage.result_size(extended) if n is None else n,
                                     stage.N, C.byref(out))
    return rc, out.value


@pytest.mark.gpu
@pytest.mark.parametrize("k", [21, 55])
def test_noalloc_checksum_equals_the_existing_entry_point_twice_in_a_row(k):
    reads = K.generate(None, 30000, seed=77 + k, n_genomes=3, genome_len_min=40000, genome_len_max=60000)
    st = K.KcountStage(k, keep_ext_counts=True)
    st.add_reads(reads)
    st.run()
    assert st.result_size(False) > 1000
    for extended in (False, True):
        want = st.checksum(extended)
        for _ in range(2):  # the second call finds the first call's partial sums in the scratch array
            rc, got = _noalloc(st, extended)
            assert rc == 0 and got == want
    # a smaller call after a larger one: stale words of the larger call must not be summed
    other = K.KcountStage(k)
    other.add_reads(reads.slice(0, 500))
    other.run()
    rc, got = _noalloc(other, False)
    assert rc == 0 and got == other.checksum(False)
    rc, got = _noalloc(st, False)
    assert rc == 0 and got == st.checksum(False)
    # empty input: checksum 0
    rc, got = _noalloc(st, False, n=0)
    assert rc == 0 and got == 0


@pytest.mark.gpu
def test_noalloc_checksum_rejects_sizes_beyond_its_maximum():
    L = _lib.lib()
    out = C.c_uint64()
    assert L.kc_checksum_noalloc(0, 1, 1, 1, 1, None, (1 << 31) + 1, 1, C.byref(out)) == _lib.KC_ERR_ARG
    assert L.kc_checksum_noalloc(0, 1, 1, 1, 1, None, 10, 5, C.byref(out)) == _lib.KC_ERR_ARG
    assert L.kc_checksum_noalloc(0, None, None, None, None, None, 10, 1, C.byref(out)) == _lib.KC_ERR_ARG
