mkdir -p gpurun_out
( timeout 200 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -3 gpurun_out/r02_noalloc_alone.log )
( timeout 300 python -m pytest tests/test_merge.py -x -q -m gpu > gpurun_out/r02_merge_gpu.log 2>&1; echo "merge rc=$?"; tail -8 gpurun_out/r02_merge_gpu.log )
timeout 200 python - <<'PY' 2>&1 | tail -5
import numpy as np, time
import mhm2_kmer_analysis_v2_b200 as K
ps = K.PairSet.from_fastq('tests/golden/merge/basic.fq.gz')
R = 1000
n = len(ps.lens)
seq = np.tile(ps.seq, R); qual = np.tile(ps.qual, R)
offs = (np.tile(ps.offsets, R).reshape(R, n) + (np.arange(R, dtype=np.uint64) * np.uint64(len(ps.seq)))[:, None]).reshape(-1)
lens = np.tile(ps.lens, R)
big = K.PairSet(seq, qual, offs, lens)
for _ in range(2):
    t = time.time(); reads, st = K.merge_pairs(big); w = time.time() - t
print('merge', len(big), 'pairs', st['merged'], 'merged; kernel ms', round(st['ms_kernel'], 2), 'pairs/s', round(len(big) / (st['ms_kernel'] / 1e3) / 1e6, 1), 'M; call wall s', round(w, 2),
      'bytes in+out GB/s', round(3 * len(seq) / (st['ms_kernel'] / 1e3) / 1e9, 1))
PY
( timeout 700 python -m pytest tests -x -q -m gpu > gpurun_out/r02_gputest_o.log 2>&1; echo "full pytest rc=$?"; tail -4 gpurun_out/r02_gputest_o.log )
