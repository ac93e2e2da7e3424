"""Wall-clock time of every C-ABI call of one stage pass (host + device), to find host-side gaps between kernels."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mhm2_kmer_analysis_v2_b200 as K
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 33
reads = K.generate("C2", n)
st = K.KcountStage(k, est_kmers=n * (150 - k + 1))
st.add_reads(reads); st.sync()
for it in range(4):
    t = [time.perf_counter()]
    st.reset(); t.append(time.perf_counter())
    st.parse(); t.append(time.perf_counter())
    st.insert(); t.append(time.perf_counter())
    s = st.finalize(); t.append(time.perf_counter())
    print("reset %.2f parse %.2f insert(call) %.2f finalize(call) %.2f total %.2f ms | dev: parse %.2f (kernel %.2f) insert %.2f" % (
        *[(t[i + 1] - t[i]) * 1e3 for i in range(4)], (t[4] - t[0]) * 1e3, s["ms_parse"], s["ms_parse_kernel"], s["ms_insert"]), flush=True)
