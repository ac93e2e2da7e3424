"""GPU, two PROCESSES on one device: the default multi-GPU data path -- exchange_supermers_pull -> kc_exchange_export /
kc_exchange_import (CUDA IPC works between processes that share a device) -> the refine / count kernels reading the
peer's bins through the mapped pointer -> kc_finalize -> release_peers -- against the oracle run with n_ranks = 2.
Per-rank dump D2 (all entries with count >= 2, the 8 extension counters, chosen extensions) must equal
Oracle(n_ranks=2).result(True, rank=r) bit for bit. Collectives run over gloo (NCCL refuses two ranks on one device);
the data path is the same one bench.py's --exchange pull uses. Replaces KmerDHT::add_supermer / flush_updates
(src/kcount/kmer_dht.cpp:248-258); owner rule src/kcount/kmer_dht.cpp:193-196."""
import datetime
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SMALL = dict(n_genomes=3, genome_len_min=30_000, genome_len_max=60_000, abundance_sigma=1.0, sub_rate=0.005)

pytestmark = pytest.mark.gpu


def _reads(K, k, n):
    return K.generate(None, n, seed=700 + k, **SMALL)


def _worker(rank, world, port, k, local_bins, m, mode, n_reads, steps, tmp):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    import mhm2_kmer_analysis_v2_b200 as K
    from mhm2_kmer_analysis_v2_b200 import exchange as X

    torch.cuda.set_device(0)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            timeout=datetime.timedelta(seconds=120))
    try:
        reads = _reads(K, k, n_reads)
        n = len(reads)
        st = K.KcountStage(k, m=m, rank=rank, n_ranks=world, device=0, keep_ext_counts=True, local_bins=local_bins,
                           est_kmers=reads.counted_kmers(k) // world)
        X.set_exchange_mode(st, mode)
        st.add_reads(reads.slice(n * rank // world, n * (rank + 1) // world))
        for step in range(steps):  # more than one step: the mappings are cached, the buffers reused
            st.reset()
            stats, xb = X.count_step_pull(st)
        res = st.result(True)
        np.savez(os.path.join(tmp, f"rank{rank}.npz"), keys=res.keys, counts=res.counts, left=res.left.view(np.uint8),
                 right=res.right.view(np.uint8), ext8=res.ext8, checksum=np.uint64(st.checksum(True)), pulled=xb[1],
                 inserted=stats["inserted_kmers"], bins=stats["local_bins"])
        st.close()
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


# (k, local_bins, m, mode): local_bins 0 = automatic (few bins at this size: the count kernel reads both sources'
# segments itself); >= 65536 = two-level partition (the refine kernel reads the peer's coarse bins and merges them)
CASES = [(21, 0, 0, "pull"), (21, 100_000, 0, "pull"), (55, 0, 0, "pull"), (55, 131_072, 0, "pull"), (99, 0, 0, "pull"),
         (99, 70_000, 15, "pull"), (33, 100_000, 0, "copy"), (77, 80_000, 15, "copy"), (21, 0, 0, "copy")]


@pytest.mark.parametrize("k,local_bins,m,mode", CASES)
def test_two_processes_pull_matches_oracle(oracle_mod, tmp_path, k, local_bins, m, mode):
    import mhm2_kmer_analysis_v2_b200 as K

    world, n_reads = 2, 8_000
    mp.spawn(_worker, args=(world, _free_port(), k, local_bins, m, mode, n_reads, 2, str(tmp_path)), nprocs=world, join=True)
    reads = _reads(K, k, n_reads)
    o = oracle_mod.Oracle(k, m=m, n_ranks=world, slots_per_rank=4 * reads.counted_kmers(k))
    o.add_reads(reads.bytes, reads.offsets, reads.lens)
    o.finalize()
    total_inserted = 0
    for r in range(world):
        p = np.load(os.path.join(str(tmp_path), f"rank{r}.npz"))
        ref = o.result(True, rank=r)
        assert len(p["counts"]) == len(ref), f"rank {r}: {len(p['counts'])} entries vs oracle {len(ref)}"
        assert np.array_equal(p["keys"], ref.keys), f"rank {r}: keys"
        for name, want in (("counts", ref.counts), ("left", ref.left.view(np.uint8)), ("right", ref.right.view(np.uint8)), ("ext8", ref.ext8)):
            assert np.array_equal(p[name], want), f"rank {r}: {name}"
        assert int(p["checksum"]) == o.checksum(True, rank=r)
        assert int(p["pulled"]) > 0, "nothing came from the peer"
        if local_bins >= 65536:
            assert int(p["bins"]) >= 65536
        total_inserted += int(p["inserted"])
    assert total_inserted == reads.counted_kmers(k)
