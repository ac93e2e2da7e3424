#!/usr/bin/env python
"""Generates the golden vectors of tests/golden/*.npz by running the REFERENCE's own CPU kcount
(oracle/_ref/kcount_ref = /root/reference/src/{kmer.cpp,hash_funcs.c,kcount/kcount.cpp,kcount/kcount_cpu.cpp,kcount/kmer_dht.cpp}
compiled unmodified behind oracle/ref_shim) on small seeded read sets. Only runs where /root/reference exists:

    make -C oracle ref && python tests/golden/make_golden.py

The reference ships no tests or golden vectors for this path, so "outputs of the reference itself
run here" is the pin. Each .npz holds the reads (PackedRead bytes / offsets / lens, or a recipe when they are large), the
parameters, the reference's JSON summary and its sorted dumps D1 (survivors) and D2 (all entries with count >= 2 with the
8 extension counters and the chosen extensions incl. X / F).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import mhm2_kmer_analysis_v2_b200 as K  # noqa: E402
import oracle as O  # noqa: E402

TMP = "/tmp/kc_golden"


def edge_reads(k):
    """Hand-made reads around the quirks listed in SURVEY.md 7.3 / 8c."""
    rng = np.random.default_rng(k)
    g = "".join(rng.choice(list("ACGT"), 400))
    seqs, quals = [], []

    def add(s, q=None):
        seqs.append(s)
        quals.append(q if q is not None else "I" * len(s))

    for _ in range(3):  # a few clean copies so that count >= 2
        add(g[:200])
        add(g[100:300])
    add(g[:k - 1])            # shorter than k: skipped (kcount.cpp:78)
    add(g[:k])                # one k-mer, never counted
    add(g[:k + 1])            # two k-mers, none interior
    add(g[:k + 2])            # exactly one counted k-mer
    add(g[:k + 2])
    add(g[50:50 + k + 2], "I" + "I" * k + "+")   # low-quality right extension
    add(g[50:50 + k + 2], "+" + "I" * k + "I")   # low-quality left extension
    s = list(g[20:180])
    s[40] = "N"               # N inside k-mers (-> G) and as an extension (never counted)
    s[41] = "N"
    add("".join(s))
    add("".join(s))
    add("N")                  # merged-pair placeholder read (merge_reads.cpp:376)
    add("")                   # empty read
    lowq = "".join("+" if i % 7 == 0 else "I" for i in range(200))
    add(g[:200], lowq)
    rc = g[:200][::-1].translate(str.maketrans("ACGT", "TGCA"))
    add(rc)                   # reverse strand: canonicalisation + extension swap/complement
    add(rc)
    if k % 2 == 0:            # palindromic k-mers exist only for even k
        half = g[300:300 + k // 2]
        pal = half + half[::-1].translate(str.maketrans("ACGT", "TGCA"))
        for _ in range(3):
            add("AC" + pal + "GT")
            add("TT" + pal + "CA")
    add("A" * 150)            # homopolymer: every k-mer identical, extensions A/A
    add("A" * 150)
    add("ACGT" * 37)          # period-4 repeat
    return K.ReadSet.from_strings(seqs, quals)


def synth(seed, n, **kw):
    base = dict(n_genomes=2, genome_len_min=3000, genome_len_max=6000, abundance_sigma=0.7, sub_rate=0.01, read_len=150,
                read_len_min=40, n_rate=0.002)
    base.update(kw)
    return K.generate(None, n, seed=seed, **base)


def saturation_reads():
    """70 000 copies of one read (+ its reverse complement 5 times): count and extension counters saturate at 65535."""
    rng = np.random.default_rng(99)
    s = "".join(rng.choice(list("ACGT"), 60))
    one = K.ReadSet.from_strings([s])
    n = 70_000
    bytes_ = np.tile(one.bytes, n)
    offs = (np.arange(n, dtype=np.uint64) * np.uint64(len(s)))
    lens = np.full(n, len(s), np.uint16)
    return K.ReadSet(bytes_, offs, lens), dict(recipe="saturation", seq=s, copies=n)


CASES = []
for k in (21, 33, 55, 77, 99):
    CASES.append((f"synth_k{k}", k, 2, lambda k=k: synth(7000 + k, 700)))
    CASES.append((f"edge_k{k}", k, 2, lambda k=k: edge_reads(k)))
CASES.append(("edge_k22_palindromes", 22, 2, lambda: edge_reads(22)))
CASES.append(("edge_k34_palindromes", 34, 2, lambda: edge_reads(34)))
CASES.append(("synth_k31_dmin3", 31, 3, lambda: synth(7031, 700)))
CASES.append(("deep_k21_dyn_min_depth", 21, 2, lambda: synth(7121, 4000, n_genomes=1, genome_len_min=600, genome_len_max=600,
                                                             abundance_sigma=0.0, sub_rate=0.03, read_len_min=0, n_rate=0.0)))


def run_case(name, k, dmin, reads, recipe=None):
    os.makedirs(TMP, exist_ok=True)
    rp = os.path.join(TMP, name + ".reads")
    reads.save(rp)
    s = O.run_reference(rp, k, os.path.join(TMP, name), extended=True, dmin=dmin)
    assert s["dropped"] in (0, -1), s
    d1, d2 = O.load_dump(os.path.join(TMP, name + ".d1")), O.load_dump(os.path.join(TMP, name + ".d2"))
    out = dict(k=k, dmin=dmin, summary=json.dumps(s), d1_keys=d1.keys, d1_counts=d1.counts, d1_left=d1.left.view(np.uint8),
               d1_right=d1.right.view(np.uint8), d2_keys=d2.keys, d2_counts=d2.counts, d2_left=d2.left.view(np.uint8),
               d2_right=d2.right.view(np.uint8), d2_ext8=d2.ext8)
    if recipe is None:
        out.update(bytes=reads.bytes, offsets=reads.offsets, lens=reads.lens)
    else:
        out.update(recipe=json.dumps(recipe))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(f"{name}: k={k} reads={len(reads)} entries={s['entries']} ge2={len(d2)} final={len(d1)} max_count={s['max_count']}")


if __name__ == "__main__":
    if not os.path.exists(O.REF_BIN):
        raise SystemExit("oracle/_ref/kcount_ref missing: run `make -C oracle ref` where /root/reference exists")
    for name, k, dmin, make in CASES:
        run_case(name, k, dmin, make())
    r, recipe = saturation_reads()
    run_case("saturation_k21", 21, 2, r, recipe)
