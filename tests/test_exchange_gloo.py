"""CPU, world_size 2 and 3 over gloo: the host side of the owner exchange (mhm2_kmer_analysis_v2_b200/exchange.py) --
size exchange, receive layout, all-to-all, stats reductions -- driven with a stand-in stage whose supermers come from the
oracle's process_seq restatement. The union of the owners' tables must equal the single-rank oracle result, which is
what replaces the reference's add_supermer RPC + flush_updates + barrier (src/kcount/kmer_dht.cpp:248-258)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODES = "ACGTN"


class OracleStage:
    """Host stand-in with the interface exchange_supermers() needs. Wire format per supermer: u16 length + chars
    (upper case = high quality, as the reference's CPU build sends them, kcount_cpu.cpp:86-101)."""

    def __init__(self, O, k, rank, n_ranks, reads):
        self.O, self.k, self.rank, self.n_ranks = O, k, rank, n_ranks
        self.oracle = O.Oracle(k, n_ranks=n_ranks, slots_per_rank=max(4 * reads.counted_kmers(k), 1 << 12))
        out = [bytearray() for _ in range(n_ranks)]
        for i in range(len(reads)):
            b = reads.bytes[int(reads.offsets[i]): int(reads.offsets[i]) + int(reads.lens[i])]
            for target, first, n in self.oracle.read_supermers(b):
                seg = b[first - 1: first + n + k]
                s = "".join(CODES[min(int(x) & 7, 4)] if (int(x) >> 3) >= 20 else CODES[min(int(x) & 7, 4)].lower() for x in seg)
                out[target] += len(s).to_bytes(2, "little") + s.encode()
        self.send = [torch.tensor(np.frombuffer(bytes(o), np.uint8).copy()) for o in out]
        self.recv = None

    def send_sizes(self):
        return np.array([len(t) for t in self.send], np.uint64)

    def send_view(self, d):
        return self.send[d] if d != self.rank else torch.empty(0, dtype=torch.uint8)

    def recv_buffer(self, recv_sizes):
        self.recv_sizes = [int(x) for x in recv_sizes]
        self.recv = [torch.empty(n, dtype=torch.uint8) for n in self.recv_sizes]
        off = np.concatenate([[0], np.cumsum(self.recv_sizes)[:-1]]).astype(np.uint64)
        return 0, off

    def recv_view(self, s):
        return self.recv[s]

    def exchange_done(self):
        self.recv[self.rank] = self.send[self.rank]

    def insert_and_finalize(self):
        n_sm = 0
        for buf in self.recv:
            raw = buf.numpy().tobytes()
            at = 0
            while at < len(raw):
                ln = int.from_bytes(raw[at: at + 2], "little")
                self.oracle.insert_supermer(self.rank, raw[at + 2: at + 2 + ln].decode())
                at += 2 + ln
                n_sm += 1
        self.oracle.finalize()
        return n_sm


def _worker(rank, world, port, k, tmp):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import mhm2_kmer_analysis_v2_b200 as K
    import oracle as O
    from mhm2_kmer_analysis_v2_b200.exchange import exchange_supermers, global_avg_kmer_count, reduce_stats

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        reads = K.generate(None, 900, seed=31 + k, n_genomes=2, genome_len_min=4000, genome_len_max=6000, read_len_min=60, n_rate=0.002)
        n = len(reads)
        mine = reads.slice(n * rank // world, n * (rank + 1) // world)
        st = OracleStage(O, k, rank, world, mine)
        sent, received, secs = exchange_supermers(st)
        assert sent == int(st.send_sizes().sum()) - len(st.send[rank]) and secs >= 0
        n_sm = st.insert_and_finalize()
        res = st.oracle.result(True, rank=rank)
        fin = st.oracle.result(False, rank=rank)
        stats = dict(n_reads=len(mine), n_bases=mine.n_bases, raw_kmers=0, counted_kmers=mine.counted_kmers(k), supermers_out=0,
                     supermer_bytes_out=sent, supermers_in=n_sm, inserted_kmers=0, unique_kmers=st.oracle.num_entries(rank),
                     ge2_kmers=len(res), final_kmers=len(fin), purged_kmers=0, sum_counts=int(fin.counts.astype(np.int64).sum()),
                     table_bytes=0, max_count=int(res.counts.max()) if len(res) else 0, ms_parse=0.0, ms_insert=0.0, ms_finalize=float(rank))
        tot = reduce_stats(stats)
        avg = global_avg_kmer_count(stats)
        np.savez(os.path.join(tmp, f"rank{rank}.npz"), keys=res.keys, counts=res.counts, left=res.left.view(np.uint8),
                 right=res.right.view(np.uint8), ext8=res.ext8, tot_final=tot["final_kmers"], tot_reads=tot["n_reads"],
                 tot_sum=tot["sum_counts"], avg=avg, ms_fin=tot["ms_finalize"], received=received)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,k", [(2, 21), (2, 55), (3, 33)])
def test_exchange_union_equals_single_rank(oracle_mod, tmp_path, world, k):
    import mhm2_kmer_analysis_v2_b200 as K

    mp.spawn(_worker, args=(world, _free_port(), k, str(tmp_path)), nprocs=world, join=True)
    reads = K.generate(None, 900, seed=31 + k, n_genomes=2, genome_len_min=4000, genome_len_max=6000, read_len_min=60, n_rate=0.002)
    o = oracle_mod.Oracle(k, slots_per_rank=4 * reads.counted_kmers(k))
    o.add_reads(reads.bytes, reads.offsets, reads.lens)
    o.finalize()
    ref, fin = o.result(True), o.result(False)
    parts = [np.load(os.path.join(str(tmp_path), f"rank{r}.npz")) for r in range(world)]
    keys = np.concatenate([p["keys"] for p in parts])
    order = np.lexsort([keys[:, i] for i in reversed(range(keys.shape[1]))])
    assert np.array_equal(keys[order], ref.keys)
    for name, want in (("counts", ref.counts), ("left", ref.left.view(np.uint8)), ("right", ref.right.view(np.uint8)), ("ext8", ref.ext8)):
        assert np.array_equal(np.concatenate([p[name] for p in parts])[order], want), name
    for p in parts:  # every rank holds the same reduced totals
        assert int(p["tot_final"]) == len(fin) and int(p["tot_reads"]) == len(reads)
        assert int(p["tot_sum"]) == int(fin.counts.astype(np.int64).sum())
        assert abs(float(p["avg"]) - o.avg_kmer_count()) < 1e-9
        assert float(p["ms_fin"]) == world - 1  # MAX reduction
    assert sum(int(p["received"]) for p in parts) > 0
