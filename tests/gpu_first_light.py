"""Manual bring-up script (not collected by pytest): run on the GPU box, prints diagnostics instead of asserting."""
import sys, os, time, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
import mhm2_kmer_analysis_v2_b200 as K
import oracle as O
from util import oracle_run, gpu_single

SMALL = dict(n_genomes=3, genome_len_min=30_000, genome_len_max=60_000, abundance_sigma=1.0, sub_rate=0.005)
for k in (21, 33, 55, 77, 99):
    try:
        reads = K.generate(None, 6000, seed=100 + k, **SMALL)
        o = oracle_run(O, reads, k)
        st, stats = gpu_single(reads, k)
        r, ref = st.result(True), o.result(True)
        same = len(r) == len(ref) and np.array_equal(r.keys, ref.keys) and np.array_equal(r.counts, ref.counts) \
            and np.array_equal(r.ext8, ref.ext8) and np.array_equal(r.left, ref.left) and np.array_equal(r.right, ref.right)
        print(f"k={k} same={same} n={len(r)}/{len(ref)} unique={stats['unique_kmers']}/{o.num_entries()} "
              f"occ={stats['inserted_kmers']}/{o.counted_kmers()} ms parse={stats['ms_parse']:.3f} insert={stats['ms_insert']:.3f} "
              f"fin={stats['ms_finalize']:.3f}", flush=True)
        if not same:
            kset = {tuple(x) for x in r.keys.tolist()}
            rset = {tuple(x) for x in ref.keys.tolist()}
            print("  missing", len(rset - kset), "extra", len(kset - rset))
            if len(r) == len(ref) and np.array_equal(r.keys, ref.keys):
                for nm in ("counts", "ext8", "left", "right"):
                    a, b = getattr(r, nm), getattr(ref, nm)
                    d = np.nonzero((a != b).reshape(len(r), -1).any(axis=1))[0]
                    print("  ", nm, "diffs", len(d), (a[d[0]], b[d[0]]) if len(d) else "")
        st.close()
    except Exception:
        traceback.print_exc()
# a first throughput look: C1-like 1M reads, k=21
reads = K.generate("C1", 1_000_000)
for bins in (1, 64, 1024):
    try:
        st = K.KcountStage(21, local_bins=bins)
        st.add_reads(reads)
        for it in range(3):
            st.reset()
            s = st.run()
        print(f"C1 1M reads k=21 bins={bins}: parse {s['ms_parse']:.2f} ms insert {s['ms_insert']:.2f} ms finalize {s['ms_finalize']:.2f} ms "
              f"-> {s['counted_kmers'] / (s['ms_parse'] + s['ms_insert'] + s['ms_finalize']) / 1e6:.2f} G kmers/s; unique {s['unique_kmers']} final {s['final_kmers']}", flush=True)
        st.close()
    except Exception:
        traceback.print_exc()
