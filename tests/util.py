"""Shared helpers of the parity tests: run the oracle and the CUDA stage on the same reads and compare everything."""
import numpy as np

import mhm2_kmer_analysis_v2_b200 as K


def oracle_run(O, reads, k, n_ranks=1, **kw):
    slots = max(4 * reads.counted_kmers(k) // max(n_ranks, 1), 1 << 16)
    o = O.Oracle(k, n_ranks=n_ranks, slots_per_rank=slots, **kw)
    o.add_reads(reads.bytes, reads.offsets, reads.lens)
    o.finalize()
    return o


def assert_same(res, ref, what):
    assert len(res) == len(ref), f"{what}: {len(res)} entries vs oracle {len(ref)}"
    for name in ("keys", "counts", "left", "right", "ext8"):
        a, b = getattr(res, name), getattr(ref, name)
        if a is None or b is None:
            continue
        if not np.array_equal(a, b):
            d = np.nonzero((a != b).reshape(len(res), -1).any(axis=1))[0]
            i = int(d[0])
            raise AssertionError(f"{what}: {name} differs in {len(d)} entries, first at {i}: {a[i]} vs oracle {b[i]}")


def gpu_single(reads, k, keep_ext=True, **kw):
    st = K.KcountStage(k, keep_ext_counts=keep_ext, **kw)
    st.add_reads(reads)
    stats = st.run()
    return st, stats


def local_exchange(stages):
    """All-to-all between stages that live in this process on one GPU: device-to-device copies of the per-owner
    buffers (what NCCL does between processes). Exercises kc_exchange_* without a second GPU."""
    import torch

    from mhm2_kmer_analysis_v2_b200.exchange import device_bytes

    G = len(stages)
    dev = torch.device("cuda", torch.cuda.current_device())
    sizes = np.stack([np.asarray(s.send_sizes(), np.uint64) for s in stages])  # [src, dst]
    for d, st in enumerate(stages):
        recv_sizes = sizes[:, d].copy()
        recv_sizes[d] = 0
        ptr, off = st.recv_buffer(recv_sizes)
        for s in range(G):
            n = int(recv_sizes[s])
            if s == d or n == 0:
                continue
            stages[s].sync()
            src = device_bytes(stages[s].send_ptr(d), n, dev)
            dst = device_bytes(ptr + int(off[s]), n, dev)
            dst.copy_(src)
        torch.cuda.synchronize()
        st.exchange_done()
    return sizes


def gpu_multi(reads, k, G, keep_ext=True, **kw):
    """G ranks emulated on one GPU: rank r parses a contiguous 1/G slice of the reads."""
    n = len(reads)
    stages = []
    for r in range(G):
        st = K.KcountStage(k, rank=r, n_ranks=G, keep_ext_counts=keep_ext, **kw)
        st.add_reads(reads.slice(n * r // G, n * (r + 1) // G))
        st.parse()
        stages.append(st)
    local_exchange(stages)
    stats = []
    for st in stages:
        st.insert()
        stats.append(st.finalize())
    return stages, stats


def merge_results(results, N):
    """Union of per-rank results (disjoint key sets), sorted by key words."""
    keys = np.concatenate([r.keys for r in results]).reshape(-1, N)
    order = np.lexsort([keys[:, i] for i in reversed(range(N))])
    ext = None
    if all(r.ext8 is not None for r in results):
        ext = np.concatenate([r.ext8 for r in results])[order]
    return K.KmerResult(keys[order], np.concatenate([r.counts for r in results])[order],
                        np.concatenate([r.left for r in results])[order], np.concatenate([r.right for r in results])[order], ext)
