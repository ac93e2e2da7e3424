"""GPU: the REFERENCE's own analyze_kmers / count_kmers / KmerDHT (src/kcount/kcount.cpp, src/kcount/kmer_dht.cpp, src/kmer.cpp
compiled unmodified) running over the B200 binding -- integration/kcount_b200_glue.cpp in place of kcount_cpu.cpp at the
link-time plug point of src/kcount/CMakeLists.txt:24-69 (SeqBlockInserter<MAX_K>, src/kcount/kcount.hpp:56-68;
HashTableInserter<MAX_K>, src/kcount/kmer_dht.hpp:95-116), linked against libkcount_b200.so: oracle/_ref/kcount_ref_b200
(`make -C oracle ref_b200`, built where /root/reference exists, travels to the GPU box like the other built files).

Its dump D1 -- read back through the reference's KmerDHT::local_kmers_begin / local_kmers_end
(src/kcount/kmer_dht.cpp:300-307) -- must equal the CPU build's (oracle/_ref/kcount_ref -> tests/golden/*.npz) bit for
bit, with one rank and with two forked ranks sharing the device (owner exchange over CUDA IPC from C++)."""
import json
import os
import subprocess

import numpy as np
import pytest

from test_oracle_golden import GOLDEN, load_golden

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "kcount_ref_b200")
NAMES = ["synth_k21", "synth_k33", "synth_k55", "synth_k77", "synth_k99", "edge_k21", "edge_k33", "edge_k99", "edge_k34_palindromes",
         "synth_k31_dmin3", "deep_k21_dyn_min_depth", "saturation_k21"]


def _run(tmp_path, name, ranks):
    import oracle as O

    if not os.path.exists(BIN):
        pytest.fail("oracle/_ref/kcount_ref_b200 is missing: run `make -C oracle ref_b200` where /root/reference exists")
    path = [p for p in GOLDEN if os.path.basename(p) == name + ".npz"][0]
    d, k, dmin, reads, summary = load_golden(path)
    rp = str(tmp_path / "reads.bin")
    reads.save(rp)
    pre = str(tmp_path / "out")
    r = subprocess.run([BIN, "--reads", rp, "--k", str(k), "--dmin", str(dmin), "--ranks", str(ranks), "--out", pre], capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "DIE" not in r.stderr and "DIE" not in r.stdout, (r.stdout + r.stderr)[-2000:]
    s = json.loads(r.stdout.strip().splitlines()[-1])
    got = O.load_dump(pre + ".d1")
    return d, summary, s, got


@pytest.mark.parametrize("name", NAMES)
def test_reference_stage_over_the_b200_binding(tmp_path, name):
    d, summary, s, got = _run(tmp_path, name, 1)
    assert len(got) == len(d["d1_counts"]), f"{len(got)} survivors vs the CPU build's {len(d['d1_counts'])}"
    assert np.array_equal(got.keys, d["d1_keys"]) and np.array_equal(got.counts, d["d1_counts"])
    assert np.array_equal(got.left.view(np.uint8), d["d1_left"]) and np.array_equal(got.right.view(np.uint8), d["d1_right"])
    assert s["final_kmers"] == summary["final_kmers"] and s["sum_count"] == summary["sum_count"]
    assert s["occurrences"] == summary["occurrences"]
    assert abs(s["avg_kmer_count"] - summary["avg_kmer_count"]) < 1e-5  # KmerDHT::get_avg_kmer_count


@pytest.mark.parametrize("name", ["synth_k21", "synth_k55", "synth_k99", "edge_k33"])
def test_two_forked_ranks_exchange_over_ipc(tmp_path, name):
    d, summary, s, got = _run(tmp_path, name, 2)
    assert s["ranks"] == 2
    assert np.array_equal(got.keys, d["d1_keys"]) and np.array_equal(got.counts, d["d1_counts"])
    assert np.array_equal(got.left.view(np.uint8), d["d1_left"]) and np.array_equal(got.right.view(np.uint8), d["d1_right"])
    assert s["sum_count"] == summary["sum_count"]
