"""The C++ façade (include/kcount_b200.hpp: the reference's analyze_kmers / KmerDHT / PackedReads names over the C ABI).
CPU: it compiles with plain g++ -std=c++17 against the header + libkcount_b200.so and reports KC_ERR_CUDA without a GPU.
GPU: the `KMER count L R` dump of a reference-shaped caller equals the oracle's survivors line by line."""
import json
import os
import subprocess

import numpy as np
import pytest

import mhm2_kmer_analysis_v2_b200 as K

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def facade_bin(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("facade") / "facade_main")
    pkg = os.path.join(ROOT, "mhm2_kmer_analysis_v2_b200")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "cpp", "facade_main.cpp"), "-L", pkg, "-lkcount_b200", "-lkc_synth", f"-Wl,-rpath,{pkg}", "-o", out],
                   check=True)
    return out


def test_facade_compiles_and_has_no_cpu_path(facade_bin, tmp_path):
    from mhm2_kmer_analysis_v2_b200 import _lib

    if _lib.lib().kc_device_count() > 0:
        pytest.skip("a CUDA device is present")
    r = subprocess.run([facade_bin, "33", "200", "5", str(tmp_path / "d.txt")], capture_output=True, text=True)
    assert r.returncode == 42 and "no CPU path" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("k", [21, 33, 77])
def test_facade_dump_equals_oracle(facade_bin, oracle_mod, tmp_path, k):
    n, seed = 6000, 900 + k
    dump = tmp_path / f"kmers-{k}.txt"
    r = subprocess.run([facade_bin, str(k), str(n), str(seed), str(dump)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    info = json.loads(r.stdout.strip().splitlines()[-1])
    reads = K.generate(None, n, seed=seed, n_genomes=3, genome_len_min=20000, genome_len_max=40000, abundance_sigma=1.0, n_rate=0.001)
    o = oracle_mod.Oracle(k, slots_per_rank=4 * reads.counted_kmers(k))
    o.add_reads(reads.bytes, reads.offsets, reads.lens)
    o.finalize()
    ref = o.result(False)
    lines = dump.read_text().splitlines()
    assert len(lines) == len(ref) == info["kmers"]
    want = [f"{s} {c} {l.decode()} {r_.decode()}" for s, c, l, r_ in zip(K.KmerResult(ref.keys, ref.counts, ref.left, ref.right).kmer_strings(k),
                                                                     ref.counts, ref.left, ref.right)]
    assert lines == want
    assert abs(info["avg_kmer_count"] - o.avg_kmer_count()) < 1e-6 and info["m"] == o.m
