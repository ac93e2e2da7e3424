#!/usr/bin/env python
"""Writes tests/golden/merge/*: seeded FASTQ fixtures for read-pair merging and what the REFERENCE makes of them.

The expected outputs come from oracle/_ref/merge_ref = the reference's own src/merge_reads.cpp, src/fastq.cpp,
src/packed_reads.cpp, src/adapters.cpp, src/utils.cpp compiled unmodified behind oracle/merge_shim (make -C oracle merge_ref;
only where /root/reference exists). Run here, commit the fixtures; the tests never need the reference tree.

Per fixture NAME: NAME.fq.gz (+ NAME_2.fq.gz for two-file input), NAME.expect.txt.gz = "seq\\nquals\\n" per read of the
reference's PackedReads after merge_reads (qualities as PackedRead keeps them: offset + min(q - offset, 31)), and an entry in
index.json {files, qual_offset, min_kmer_len, paired, reads, merged, sha256}.
"""
import gzip
import hashlib
import json
import os
import random
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, "merge")
MERGE_REF = os.path.join(ROOT, "oracle", "_ref", "merge_ref")
COMP = {"A": "T", "C": "G", "G": "C", "T": "A", "N": "N"}
QSET = (41, 40, 40, 40, 38, 35, 30, 25, 20, 12, 8)


def rc(s):
    return "".join(COMP[c] for c in reversed(s))


def genome(rng, n, period=0):
    if period:  # tandem repeats: several offsets look like an overlap -> "ambiguous"
        unit = "".join(rng.choice("ACGT") for _ in range(period))
        return (unit * (n // period + 1))[:n]
    return "".join(rng.choice("ACGT") for _ in range(n))


def mate(rng, frag, length, qo, err, n_rate, qset):
    s = list(frag[:length])
    q = [rng.choice(qset) for _ in s]
    for j in range(len(s)):
        if rng.random() < err:
            s[j] = rng.choice([c for c in "ACGT" if c != s[j]])
            q[j] = rng.choice([2, 5, 8, 12, 15, 20])
        if rng.random() < n_rate:
            s[j] = "N"
            q[j] = 2
    return "".join(s), "".join(chr(qo + x) for x in q)


def pick(rng, v):
    return v if isinstance(v, int) else rng.randint(*v)


def pairs(seed, n, insert=(230, 60), lens=(150, 150), err=0.01, n_rate=0.002, qo=33, period=0, qset=QSET):
    """n fragments -> list of (seq1, q1, seq2, q2)"""
    rng = random.Random(seed)
    g = genome(rng, 60000, period)
    out = []
    for _ in range(n):
        l1, l2 = pick(rng, lens[0]), pick(rng, lens[1])
        ins = max(8, int(rng.gauss(*insert)))
        st = rng.randrange(0, len(g) - ins)
        frag = g[st:st + ins]
        s1, q1 = mate(rng, frag, l1, qo, err, n_rate, qset)
        s2, q2 = mate(rng, rc(frag), l2, qo, err, n_rate, qset)
        out.append((s1, q1, s2, q2))
    return out


def name_of(i, which, style):
    if style == "slash":
        return f"@frag{i}/{which}"
    if style == "illumina":  # "@name 1:N:0:ACGT" -> name/1 (src/fastq.cpp:117-124)
        return f"@M01:23:FC:1:1101:{1000 + i}:2000 {which}:N:0:ACGT"
    if style == "hudson":  # "@name-R1" -> name/1 (src/fastq.cpp:94-101)
        return f"@frag{i}-R{which}"
    if style == "slash_comment":  # "@name/1 comment" (src/fastq.cpp:111-115)
        return f"@frag{i}/{which} extra words"
    raise ValueError(style)


def write_fastq(path, records):
    with gzip.GzipFile(path, "wb", mtime=0) as f:
        for name, s, q in records:
            f.write(f"{name}\n{s}\n+\n{q}\n".encode())


def fixture(name, ps, style="slash", qo=33, min_k=21, layout="interleaved", drop=()):
    """layout: interleaved | two_files | unpaired. drop: (pair index, mate) records left out of an interleaved file."""
    recs1, recs2 = [], []
    for i, (s1, q1, s2, q2) in enumerate(ps):
        r1, r2 = (name_of(i, 1, style), s1, q1), (name_of(i, 2, style), s2, q2)
        if layout == "two_files":
            recs1.append(r1)
            recs2.append(r2)
        elif layout == "unpaired":
            recs1.append((f"@single{2 * i}", s1, q1))
            recs1.append((f"@single{2 * i + 1}", s2, q2))
        else:
            if (i, 1) not in drop:
                recs1.append(r1)
            if (i, 2) not in drop:
                recs1.append(r2)
    files = [f"{name}.fq.gz"] + ([f"{name}_2.fq.gz"] if layout == "two_files" else [])
    write_fastq(os.path.join(OUT, files[0]), recs1)
    if layout == "two_files":
        write_fastq(os.path.join(OUT, files[1]), recs2)
    # the reference reads plain files
    with tempfile.TemporaryDirectory() as td:
        plain = []
        for fn in files:
            p = os.path.join(td, fn[:-3])
            with gzip.open(os.path.join(OUT, fn), "rb") as g, open(p, "wb") as o:
                o.write(g.read())
            plain.append(p)
        spec = {"interleaved": plain[0], "unpaired": plain[0] + ":", "two_files": ":".join(plain)}[layout]
        out_txt = os.path.join(td, "out.txt")
        subprocess.run([MERGE_REF, str(qo), str(min_k), out_txt, spec], check=True, cwd=td, stderr=subprocess.DEVNULL)
        lines = open(out_txt).read().split("\n")
    assert lines[0].startswith("#avg_read_len")
    body = lines[1:]
    if body and body[-1] == "":
        body.pop()
    assert len(body) % 3 == 0
    reads = [(body[i + 1], body[i + 2]) for i in range(0, len(body), 3)]
    text = "".join(f"{s}\n{q}\n" for s, q in reads).encode()
    with gzip.GzipFile(os.path.join(OUT, f"{name}.expect.txt.gz"), "wb", mtime=0) as f:
        f.write(text)
    merged = sum(1 for i in range(0, len(reads), 2) if reads[i + 1][0] == "N" and len(reads[i + 1][0]) == 1 and layout != "unpaired")
    return {"files": files, "layout": layout, "qual_offset": qo, "min_kmer_len": min_k, "paired": layout != "unpaired", "reads": len(reads),
            "merged_or_dummy_mates": merged, "avg_read_len": int(lines[0].split()[1]), "sha256": hashlib.sha256(text).hexdigest()}


def main():
    os.makedirs(OUT, exist_ok=True)
    idx = {}
    idx["basic"] = fixture("basic", pairs(11, 600))
    idx["long_inserts"] = fixture("long_inserts", pairs(12, 300, insert=(330, 60)))  # most pairs do not overlap
    idx["short_inserts"] = fixture("short_inserts", pairs(13, 300, insert=(90, 30)))  # mates longer than the fragment
    idx["ragged"] = fixture("ragged", pairs(14, 400, lens=((30, 250), (30, 250)), insert=(200, 80)))  # unequal lengths: start_i > 0
    idx["many_n"] = fixture("many_n", pairs(15, 300, n_rate=0.02))  # N against N, more than 3 Ns: aborts
    idx["noisy"] = fixture("noisy", pairs(16, 400, err=0.04))  # near the mismatch / perror thresholds
    idx["repeats"] = fixture("repeats", pairs(17, 300, period=7, err=0.002))  # tandem repeats: ambiguous overlaps
    idx["repeats_long"] = fixture("repeats_long", pairs(18, 300, period=31, err=0.01))
    idx["tiny"] = fixture("tiny", pairs(19, 200, lens=((8, 40), (8, 40)), insert=(40, 15)), min_k=21)  # dropped pairs, len < overlap floor
    idx["phred64"] = fixture("phred64", pairs(20, 200, qo=64), qo=64)
    idx["illumina_names"] = fixture("illumina_names", pairs(21, 150), style="illumina")
    idx["hudson_names"] = fixture("hudson_names", pairs(22, 150), style="hudson")
    idx["comment_names"] = fixture("comment_names", pairs(23, 150), style="slash_comment")
    idx["two_files"] = fixture("two_files", pairs(24, 300), layout="two_files")
    # (an unpaired "file:" spec trips an assertion of the reference's read-count estimate under the one-rank shim: not pinned here;
    #  tests/test_merge.py checks that path by its definition, src/merge_reads.cpp:352-371)
    # (files with a missing mate likewise: the reference's estimate_num_reads pass asserts before the pair loop is reached)
    idx["high_quality"] = fixture("high_quality", pairs(28, 200, qset=(41, 41, 40, 39)))  # quality sums hit the cap of 41
    json.dump(idx, open(os.path.join(OUT, "index.json"), "w"), indent=1, sort_keys=True)
    for k, v in idx.items():
        print(k, v["reads"], v["merged_or_dummy_mates"], v["sha256"][:12])


if __name__ == "__main__":
    main()
