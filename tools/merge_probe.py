"""Merges R copies of the 'basic' golden pair set (kc_merge_pairs); prints the kernel time. Used under ncu (tools/run_r.sh)."""
import sys

import numpy as np

import mhm2_kmer_analysis_v2_b200 as K

R = int(sys.argv[1]) if len(sys.argv) > 1 else 200
ps = K.PairSet.from_fastq('tests/golden/merge/basic.fq.gz')
n = len(ps.lens)
seq = np.tile(ps.seq, R)
qual = np.tile(ps.qual, R)
offs = (np.tile(ps.offsets, R).reshape(R, n) + (np.arange(R, dtype=np.uint64) * np.uint64(len(ps.seq)))[:, None]).reshape(-1)
big = K.PairSet(seq, qual, offs, np.tile(ps.lens, R))
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 1):
    reads, st = K.merge_pairs(big)
print('merge', len(big), 'pairs', st['merged'], 'merged; kernel ms', round(st['ms_kernel'], 3), 'M pairs/s', round(len(big) / (st['ms_kernel'] / 1e3) / 1e6, 1))
