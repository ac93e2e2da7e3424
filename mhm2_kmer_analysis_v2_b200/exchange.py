"""Owner exchange of supermer records: the replacement for the reference's UPC++ aggregated RPCs
(KmerDHT::add_supermer / flush_updates, src/kcount/kmer_dht.cpp:248-258; ThreeTierAggrStore, source absent).

One process per GPU. After the parse phase every rank holds, per destination owner, one contiguous device buffer
([bin header][records]); sizes are exchanged first, then one all-to-all moves the buffers HBM -> HBM over NVLink
(NCCL). The stage object only has to offer send_sizes()/send_ptr()/recv_buffer()/exchange_done(), so the same driver
runs against the CUDA stage (nccl) and against a CPU stand-in in the gloo tests.
"""
import numpy as np
import torch
import torch.distributed as dist


class _DevBytes:
    """Zero-copy torch view of raw device memory (no ownership)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def device_bytes(ptr, nbytes, device):
    if nbytes == 0:
        return torch.empty(0, dtype=torch.uint8, device=device)
    return torch.as_tensor(_DevBytes(ptr, nbytes), device=device)


def _is_cuda_stage(stage):
    return hasattr(stage, "L")  # KcountStage (C ABI) as opposed to a host stand-in used by the CPU tests


def exchange_sizes(send_sizes, group=None, device="cpu"):
    """all-to-all of one u64 per pair: recv[s] = what rank s will send me."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    # all-gather of the G x G size matrix (works on nccl and gloo alike; G*G*8 bytes)
    t_send = torch.as_tensor(np.asarray(send_sizes, np.int64), device=device)
    rows = [torch.empty(world, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(rows, t_send, group=group)
    return torch.stack(rows)[:, rank].cpu().numpy().astype(np.uint64)


def _all_to_all(recvs, sends, group=None):
    """Variable-size all-to-all. NCCL: one grouped ncclSend/ncclRecv all-to-all; gloo (CPU tests) has no alltoall, so
    the same exchange is posted as non-blocking point-to-point pairs."""
    if dist.get_backend(group) == "nccl":
        dist.all_to_all(recvs, sends, group=group)
        return
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    reqs = []
    for peer in range(world):
        if peer == rank:
            continue
        if recvs[peer].numel():
            reqs.append(dist.irecv(recvs[peer], src=dist.get_global_rank(group, peer) if group else peer, group=group))
        if sends[peer].numel():
            reqs.append(dist.isend(sends[peer], dst=dist.get_global_rank(group, peer) if group else peer, group=group))
    for r in reqs:
        r.wait()


def exchange_supermers(stage, group=None):
    """Move every rank's per-owner supermer buffers to their owners. Returns (bytes_sent, bytes_received, seconds)
    where seconds is the device time of the data all-to-all (CUDA events) on NCCL, wall time on gloo."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    assert world == stage.n_ranks and rank == stage.rank
    cuda = _is_cuda_stage(stage)
    dev = torch.device("cuda", torch.cuda.current_device()) if cuda else torch.device("cpu")
    send_sizes = np.asarray(stage.send_sizes(), np.uint64)
    send_sizes[rank] = 0
    recv_sizes = exchange_sizes(send_sizes, group, dev)
    recv_sizes[rank] = 0
    recv_ptr, recv_off = stage.recv_buffer(recv_sizes)
    if cuda:
        stage.sync()
        sends = [device_bytes(stage.send_ptr(d), int(send_sizes[d]), dev) for d in range(world)]
        recvs = [device_bytes(recv_ptr + int(recv_off[s]), int(recv_sizes[s]), dev) for s in range(world)]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _all_to_all(recvs, sends, group=group)
        e1.record()
        torch.cuda.synchronize()
        secs = e0.elapsed_time(e1) / 1e3
    else:
        import time

        sends = [stage.send_view(d) for d in range(world)]
        recvs = [stage.recv_view(s) for s in range(world)]
        t0 = time.perf_counter()
        _all_to_all(recvs, sends, group=group)
        secs = time.perf_counter() - t0
    stage.exchange_done()
    return int(send_sizes.sum()), int(recv_sizes.sum()), secs


MAX_SLICES = 12  # KC_MAX_SLICES


def set_exchange_mode(stage, mode):
    """'pull' (the kernels read the peers' memory) or 'copy' (copy engines stream it in, slice by slice, beside the
    kernels). Both are driven by exchange_supermers_pull / release_peers."""
    stage.set_exchange_mode(mode)


class ExchangeAborted(RuntimeError):
    """A rank of the group failed inside the peer-memory protocol; every rank raises this instead of waiting forever."""


def _coll_device(group):
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")


def exchange_supermers_pull(stage, group=None):
    """Peer-memory variant for the GPUs of one NVLink / NVSwitch node: nothing is moved here. Every rank publishes its
    send buffer (CUDA IPC handle + per-owner region offsets / used bytes, one small all-gather), every owner maps the
    regions meant for it, and kc_insert's refine kernel pulls the records out of the peers' HBM while it sorts them into
    fine bins -- transfer and partition in one kernel, only used records on the links. The caller must call
    release_peers() after finalize() -- also when its own insert / finalize failed (ok=False) -- before any rank parses
    again. Returns (0, bytes_to_pull, 0.0).

    Failure path: the all-gather carries a status byte. A rank whose parse / export failed still takes part (status 0)
    and every rank raises ExchangeAborted without importing anything; a rank that died is caught by the process group's
    timeout (bench.py and the tests set it), not by an endless wait."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    assert world == stage.n_ranks and rank == stage.rank
    dev = _coll_device(group)
    ok, err = 1, None
    h, off, nb = np.zeros(64, np.uint8), np.zeros(world, np.uint64), np.zeros(world, np.uint64)
    n_sl, sl = 1, np.zeros((world, MAX_SLICES + 1), np.uint64)
    try:
        stage.sync()  # this rank's bins are complete (kc_parse synchronises, this is for callers that queued more work)
        h, off, nb = stage.export_regions()
        n_sl, sl = stage.export_slices()
    except Exception as e:  # noqa: BLE001 -- reported to the peers, re-raised below
        ok, err = 0, e
    mine = torch.from_numpy(np.concatenate([np.array([ok, n_sl] + [0] * 6, np.uint8), h.view(np.uint8), off.view(np.uint8), nb.view(np.uint8),
                                            sl.reshape(-1).view(np.uint8)])).to(dev)
    rows = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(rows, mine, group=group)  # also the barrier "every rank has parsed"
    rows = [r.cpu().numpy() for r in rows]
    failed = [s for s in range(world) if rows[s][0] == 0]
    if failed:
        raise ExchangeAborted(f"peer-memory exchange aborted: rank(s) {failed} failed before the export" + (f" (here: {err})" if err else ""))
    pulled = 0
    for s in range(world):
        if s == rank:
            continue
        r = rows[s][8:]
        o = r[64:64 + 8 * world].view(np.uint64)
        b = r[64 + 8 * world:64 + 16 * world].view(np.uint64)
        t = r[64 + 16 * world:].view(np.uint64).reshape(world, MAX_SLICES + 1)
        stage.import_region(s, r[:64], int(o[rank]), int(b[rank]))
        stage.import_slices(s, int(rows[s][1]), t[rank, :int(rows[s][1]) + 1])
        pulled += int(b[rank])
    stage.exchange_done()
    return 0, pulled, 0.0


def release_peers(stage, group=None, ok=True):
    """Second half of the peer-memory exchange: my kernels are done with the peers' buffers, and theirs with mine.
    Every rank calls this once per exchange_supermers_pull, with ok=False when its insert / finalize raised: the
    agreement (an all-reduce, which is also the barrier) makes all ranks raise ExchangeAborted together instead of one
    rank waiting for a peer that will never arrive. Then the send buffers retired by a re-allocation are freed."""
    try:
        stage.sync()
    except Exception:  # noqa: BLE001
        ok = False
    t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=_coll_device(group))
    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    stage.exchange_release()
    if int(t.item()) == 0:
        raise ExchangeAborted("a rank failed between the export and the release of the peer-memory exchange")


def count_step_pull(stage, group=None):
    """parse -> peer-memory exchange -> insert -> finalize -> release, with the failure path wired: whatever this rank
    raises between export and release is reported to the peers first. Returns (stats, exchange tuple)."""
    stage.parse()
    xb = exchange_supermers_pull(stage, group)
    stats, err = None, None
    try:
        stage.insert()
        stats = stage.finalize()
    except Exception as e:  # noqa: BLE001
        err = e
    try:
        release_peers(stage, group, ok=err is None)
    except ExchangeAborted:
        if err is not None:
            raise err
        raise
    return stats, xb


def global_avg_kmer_count(stats, group=None):
    """avg depth over all ranks = sum(count) / number of k-mers (kcount_cpu.cpp:586-596, reduce_all)."""
    backend = dist.get_backend(group)
    dev = "cuda" if backend == "nccl" else "cpu"
    t = torch.tensor([float(stats["sum_counts"]), float(stats["final_kmers"])], dtype=torch.float64, device=dev)
    dist.all_reduce(t, group=group)
    return float(t[0] / t[1]) if t[1] > 0 else 0.0


def reduce_stats(stats, group=None):
    """Sum the per-rank counters the reference reduces (kcount_gpu.cpp:294-364); max for max_count / times."""
    backend = dist.get_backend(group)
    dev = "cuda" if backend == "nccl" else "cpu"
    sum_keys = ["n_reads", "n_bases", "raw_kmers", "counted_kmers", "supermers_out", "supermer_bytes_out", "supermers_in",
                "inserted_kmers", "unique_kmers", "ge2_kmers", "final_kmers", "purged_kmers", "sum_counts", "table_bytes"]
    max_keys = ["max_count", "ms_parse", "ms_insert", "ms_finalize"]
    s = torch.tensor([float(stats[k]) for k in sum_keys], dtype=torch.float64, device=dev)
    m = torch.tensor([float(stats[k]) for k in max_keys], dtype=torch.float64, device=dev)
    dist.all_reduce(s, group=group)
    dist.all_reduce(m, op=dist.ReduceOp.MAX, group=group)
    out = {k: int(v) for k, v in zip(sum_keys, s.tolist())}
    out.update({k: v for k, v in zip(max_keys, m.tolist())})
    out["max_count"] = int(out["max_count"])
    out["avg_kmer_count"] = out["sum_counts"] / out["final_kmers"] if out["final_kmers"] else 0.0
    # load balance of the owners, as the reference logs it (kcount.cpp:97-100)
    mine = torch.tensor([float(stats["inserted_kmers"])], dtype=torch.float64, device=dev)
    mx = mine.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    world = dist.get_world_size(group)
    out["owner_imbalance"] = float(mx.item() * world / max(out["inserted_kmers"], 1))
    return out
