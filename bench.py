#!/usr/bin/env python
"""Headline benchmark of the k-mer analysis stage (BASELINE.json metric: counted k-mers per second).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload = BASELINE config 2 ("C2"): seeded synthetic metagenome mix, 100 random genomes of 1-5 Mbp with lognormal(0,1)
abundance, 20 M x 150 bp reads PER GPU (weak scaling: rank r takes reads [r*20M, (r+1)*20M) of the one seeded stream),
0.5 % substitutions, k = 33 (two key words), singleton filter off, min depth 2.

One step = one complete pass of the hot path over the rank's reads: kc_reset, supermer extraction (kc_parse), the owner
exchange when N > 1 (NCCL all-to-all), table insert (kc_insert), purge/extension/compaction (kc_finalize).
  value    reads resident in HBM at the start of the step, result left in HBM; timed on the device (CUDA events on the
           stage's stream, host gaps included), max over ranks.
  e2e      same stage through the C ABI with HOST buffers: pinned reads -> kc_add_reads (H2D) -> run -> kc_result_copy of
           the final (k-mer, count, left, right) arrays into pinned host memory (D2H); wall clock between device syncs.
  roofline the dominant kernel alone (count_kernel = insert + purge + extension resolution over shared-memory tables;
           insert_kernel with --filter 1 / --global-table): B_alg(k) ALGORITHMIC bytes per counted k-mer x
           k-mers per launch / its CUDA-event time, against MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline  the reference's own CPU kcount (oracle/_ref/kcount_ref: its unmodified sources behind a one-node UPC++
           stand-in, one forked rank per host core) on a bounded prefix of the same reads; falls back to the C restatement
           (oracle/libkcount_oracle.so, one thread) when the reference build is not present.
--impl reference times only that CPU arm and prints the same JSON shape with "impl": "reference".
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

K_DEFAULT = 33
READS_PER_GPU = 20_000_000
READ_LEN = 150
WORKLOAD = "C2"
METRIC = "counted_kmers_per_sec"
UNIT = "kmers/s"


def b_alg(k, read_len=READ_LEN):
    """Algorithmic bytes per counted k-mer on the insert path (SURVEY.md 8d / BASELINE.md 3)."""
    n_words = (k + 31) // 32
    return read_len / (read_len - k - 1) + 8 * n_words + 16


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def config_dict(args, n_gpus):
    return {
        "workload": f"{WORKLOAD}: synthetic metagenome mix, 100 genomes 1-5 Mbp lognormal(0,1), {args.reads_per_gpu} x {READ_LEN} bp reads per GPU, "
                    f"0.5% substitutions, seed 2002, k={args.k}, kcount stage only",
        "k": args.k, "reads_per_gpu": args.reads_per_gpu, "read_len": READ_LEN, "n_gpus": n_gpus, "filter": bool(args.filter),
        "dmin_thres": 2, "parallelism": f"owner-partitioned x{n_gpus}" if n_gpus > 1 else "single GPU",
        "l2": "inputs (reads 3 GB, supermer records 13 GB, result 5 GB per GPU) each far larger than the 126 MB L2; no explicit flush",
    }


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.dev = device_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.dev)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def wait_first_sample(self, timeout_s=20.0):
        """Block until nvidia-smi has printed its first line: its start-up (NVML initialisation over every GPU of the box)
        holds driver locks for up to seconds and must not overlap the timed steps."""
        t0 = time.perf_counter()
        while self.proc and not self.lines and self.proc.poll() is None and time.perf_counter() - t0 < timeout_s:
            time.sleep(0.02)

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU arm
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_reference_run(k, sample_reads, n_steps, n_warmup, reads=None):
    """Times the reference CPU kcount on the first `sample_reads` reads of the workload. Returns (best-of-steps mean
    k-mers/s, dict describing the run). Uses oracle/_ref/kcount_ref when present (kind "reference"), else the oracle
    restatement (kind "port")."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O

    import mhm2_kmer_analysis_v2_b200 as K

    if reads is None:
        reads = K.generate(WORKLOAD, sample_reads)
    elif len(reads) > sample_reads:
        reads = reads.slice(0, sample_reads)
    occ = reads.counted_kmers(k)
    times = []
    if os.path.exists(O.REF_BIN):
        cores = min(host_cores(), 128)
        tmpdir = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
        path = os.path.join(tmpdir, f"kc_bench_reads_{os.getpid()}.bin")
        reads.save(path)
        try:
            for it in range(n_warmup + n_steps):
                s = O.run_reference(path, k, ranks=cores)
                if s["dropped"] not in (0, -1):
                    raise RuntimeError(f"reference run dropped {s['dropped']} k-mers")
                assert s["occurrences"] == occ
                if it >= n_warmup:
                    times.append(s["t_dht_s"] + s["t_analyze_s"])
        finally:
            os.unlink(path)
        kind = "reference"
        how = (f"reference src/kcount CPU build (oracle/_ref/kcount_ref), {cores} forked ranks over shared-memory mailboxes (no RPC cost), "
               f"KmerDHT construction + analyze_kmers")
    else:
        cores = 1
        for it in range(n_warmup + n_steps):
            t0 = time.perf_counter()
            o = O.Oracle(k, slots_per_rank=max(4 * occ, 1 << 16))
            o.add_reads(reads.bytes, reads.offsets, reads.lens)
            o.finalize()
            dt = time.perf_counter() - t0
            o.close()
            if it >= n_warmup:
                times.append(dt)
        kind = "port"
        how = "oracle/kcount_oracle.c restatement, 1 thread (reference build absent)"
    mean = sum(times) / len(times)
    info = {"value": occ / mean, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"first {len(reads)} reads of the workload ({occ} counted k-mers), {how}; {mean:.2f} s per pass"}
    return mean, occ, info


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = min(host_cores(), 128)
    sample = min(args.reads_per_gpu * args.gpus, args.cpu_sample_reads or 100_000 * cores)
    mean, occ, info = cpu_reference_run(args.k, sample, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": info["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": mean * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic", "config": config_dict(args, args.gpus), "cpu_baseline": info,
        "e2e": {"value": info["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------ GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--k", type=int, default=K_DEFAULT)
    ap.add_argument("--reads-per-gpu", type=int, default=READS_PER_GPU)
    ap.add_argument("--filter", type=int, default=0, help="1 = two-choice singleton filter on")
    ap.add_argument("--local-bins", type=int, default=0)
    ap.add_argument("--record-bytes", type=int, default=0)
    ap.add_argument("--table-slots", type=int, default=0, help="exact table size in slots (0 = the reference's sizing rule)")
    ap.add_argument("--global-table", action="store_true", help="count in one HBM table (kc_insert.cu) instead of bin-resident tables")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-sample-reads", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-lookup", action="store_true", help="skip the batched-lookup measurement against the finished result")
    ap.add_argument("--exchange", default="pull", choices=["pull", "nccl"],
                    help="multi-GPU: 'pull' = peers' bins mapped with CUDA IPC, the refine kernel reads them over NVLink; "
                         "'nccl' = all-to-all of the owner regions, then refine")
    ap.add_argument("--n-genomes", type=int, default=0, help="profiling only: fewer genomes keep the coverage of a smaller read set like the full workload's")
    ap.add_argument("--phase-detail", action="store_true", help="also print per-phase device times to stderr")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        print(f"note: warmup {args.warmup} < 3 (timing rules ask for >= 3)", file=sys.stderr)

    if args.impl == "reference":
        return run_reference_arm(args)

    import numpy as np
    import torch

    import mhm2_kmer_analysis_v2_b200 as K
    from mhm2_kmer_analysis_v2_b200.reads import PinnedBuffer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks (one process per GPU)")
        raise SystemExit(f"WORLD_SIZE {world} != --gpus {args.gpus}")
    if K._lib.lib().kc_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the stage has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from mhm2_kmer_analysis_v2_b200.exchange import exchange_supermers, exchange_supermers_pull, release_peers

    k, n = args.k, args.reads_per_gpu
    # ---- this rank's reads, generated straight into pinned host memory
    pin_reads = PinnedBuffer(n * READ_LEN)
    synth_kw = dict(K.CONFIGS[WORKLOAD]["synth"])
    if args.n_genomes:
        synth_kw["n_genomes"] = args.n_genomes
    syn = K.Synth(**synth_kw)
    t0 = time.perf_counter()
    reads = syn.generate(rank * n, n, out_bytes=pin_reads.array)
    syn.close()
    pin_off, pin_len = PinnedBuffer(n * 8), PinnedBuffer(n * 2)
    pin_off.view(np.uint64, n)[:] = reads.offsets
    pin_len.view(np.uint16, n)[:] = reads.lens
    reads = K.ReadSet.__new__(K.ReadSet)
    reads.bytes, reads.offsets, reads.lens = pin_reads.array[: n * READ_LEN], pin_off.view(np.uint64, n), pin_len.view(np.uint16, n)
    t_gen = time.perf_counter() - t0
    occ_local = reads.counted_kmers(k)
    raw_local = int(np.maximum(reads.lens.astype(np.int64) - k + 1, 0).sum())
    h2d_bytes = len(reads.bytes) + n * 10

    st = K.KcountStage(k, rank=rank, n_ranks=world, device=local_rank, est_kmers=raw_local, use_filter=bool(args.filter),
                       local_bins=args.local_bins, record_bytes=args.record_bytes, table_slots=args.table_slots,
                       global_table=args.global_table)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        st.sync()

    def one_pass():
        st.reset()
        st.parse()
        xb = (0, 0, 0.0)
        if world > 1:
            xb = exchange_supermers_pull(st) if args.exchange == "pull" else exchange_supermers(st)
        st.insert()
        fin = st.finalize()
        if world > 1 and args.exchange == "pull":
            release_peers(st)  # nobody parses again while a peer may still be reading this rank's bins
        return fin, xb

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        return float(t.item())

    # ---- device-resident arm
    st.add_reads(reads)
    st.sync()
    # the clock sampler (an nvidia-smi process) starts before the warm-up: its start-up perturbs the driver for a moment,
    # which must not fall into the timed region; the run waits for its first sample, then it samples every 200 ms through warm-up and timed steps
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        sampler.wait_first_sample()
    barrier()
    for _ in range(args.warmup):
        stats, _ = one_pass()
    launches0 = st.stats()["kernel_launches"]
    phase = {"parse": 0.0, "parse_kernel": 0.0, "insert": 0.0, "insert_kernel": 0.0, "table_clear": 0.0, "finalize": 0.0, "exchange_s": 0.0}
    xbytes = xpulled = 0
    barrier()
    st.timer_mark(0)
    step_wall = []
    for _ in range(args.steps):
        t_s = time.perf_counter()
        stats, xb = one_pass()
        step_wall.append((time.perf_counter() - t_s) * 1e3)  # (finalize returns the counters: every step ends synchronised)
        for name in ("parse", "parse_kernel", "insert", "insert_kernel", "table_clear", "finalize"):
            phase[name] += stats["ms_" + name]
        phase["exchange_s"] += xb[2]
        xbytes += xb[0]
        xpulled += xb[1]
    st.timer_mark(1)
    barrier()
    ms_total = st.timer_elapsed_ms(0, 1)
    clocks = sampler.stop() if rank == 0 else None
    launches = st.stats()["kernel_launches"] - launches0
    ms_total = max_over_ranks(ms_total)
    occ_all = sum_over_ranks(float(occ_local))
    value = occ_all * args.steps / (ms_total / 1e3)
    ms_insert_kernel = phase["insert_kernel"] / args.steps
    inserted = stats["inserted_kmers"]  # k-mers this rank's insert launch processed
    peak, peak_src = load_peaks()
    achieved = b_alg(k) * inserted / (ms_insert_kernel / 1e3) / 1e9
    traffic = None  # DRAM bytes of one count launch, from the committed ncu --set full capture (per k-mer, scaled)
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            tj = json.load(f)
        if k == K_DEFAULT and not args.global_table and not args.filter:
            traffic = tj["dram_bytes_per_counted_kmer"] * inserted
    except Exception:
        pass
    final_local = stats["final_kmers"]
    checksum_all = None
    if dist is not None:  # order-independent result checksum, summed over the owners (mod 2^64): the same for any exchange mode
        cs = st.checksum(False)
        parts = [None] * world
        dist.all_gather_object(parts, int(cs))
        checksum_all = sum(parts) % (1 << 64)

    # ---- batched lookups against the device-resident result (what the next stage asks the table)
    lookup = None
    if not args.no_lookup and final_local > 0:
        rd = st.result_device(False)
        nq = int(min(final_local, 1 << 26))
        gen = torch.Generator(device=rd["keys"].device)
        gen.manual_seed(4242)
        idx = torch.randint(0, final_local, (nq,), device=rd["keys"].device, generator=gen)
        q = rd["keys"][idx].contiguous()
        flip = 1 << (63 - 2 * ((k - 1) % 32))
        flip = flip - (1 << 64) if flip >= (1 << 63) else flip
        q[1::2, st.N - 1] ^= flip  # every second query: last base changed -> (almost always) absent
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        lc, ll, lr = st.lookup_device(q)  # first call: builds the index
        t_first = time.perf_counter() - t0
        reps = 5
        t0 = time.perf_counter()
        for _ in range(reps):
            lc, ll, lr = st.lookup_device(q)
        t_q = (time.perf_counter() - t0) / reps
        hits_ok = bool(torch.equal(lc[0::2], rd["counts"][idx[0::2]]) and torch.equal(ll[0::2], rd["left"][idx[0::2]])
                       and torch.equal(lr[0::2], rd["right"][idx[0::2]]))
        if not hits_ok:
            raise SystemExit("bench: lookup of the result's own keys returned something else")
        lookup = {"queries": nq, "queries_per_sec": nq / t_q, "ms_per_batch": t_q * 1e3, "index_build_ms": max(t_first - t_q, 0.0) * 1e3,
                  "entries": int(final_local), "hit_fraction_asked": 0.5, "miss_found": int((lc[1::2] != 0).sum().item()),
                  "where": "queries and answers resident in HBM (kc_lookup with device pointers)"}
        del q, idx, lc, ll, lr, rd

    # ---- end-to-end arm: host reads in, host result out, through the C ABI
    e2e = None
    if not args.no_e2e:
        N = st.N
        d2h_bytes = final_local * (8 * N + 4)
        pin_res = PinnedBuffer(max(d2h_bytes, 1) + 64)
        res_keys = pin_res.view(np.uint64, final_local * N)
        res_counts = pin_res.view(np.uint16, final_local, offset=final_local * N * 8)
        res_left = pin_res.view(np.uint8, final_local, offset=final_local * (N * 8 + 2))
        res_right = pin_res.view(np.uint8, final_local, offset=final_local * (N * 8 + 3))
        import ctypes as C

        def e2e_pass(detail=False):
            tt = [time.perf_counter()]

            def mark():
                if detail:
                    if detail is True:
                        st.sync()
                    tt.append(time.perf_counter())

            st.clear_reads()
            mark()
            st.add_reads(reads)            # H2D from pinned memory, asynchronous on the stage's stream
            mark()
            s, _ = one_pass()              # with the host sink set: the survivors stream out while the count kernel runs
            mark()
            assert s["final_kmers"] == final_local, "result size changed between passes"
            if detail:
                names = ["clear_reads", "add_reads(H2D)", "parse+insert+finalize(+D2H into the sink)"]
                print(f"e2e detail (ms, sync={detail}): " +
                      ", ".join(f"{n} {1e3 * (b - a):.1f}" for n, a, b in zip(names, tt, tt[1:])), file=sys.stderr)
            return s

        st.set_result_sink(res_keys, res_counts, res_left, res_right)
        for _ in range(max(2, min(args.warmup, 3))):  # warm-up: buffer pools and bin layout settle over the first two passes
            e2e_pass()
        if args.phase_detail:
            e2e_pass(True)
        barrier()
        t0 = time.perf_counter()
        e2e_wall = []
        for _ in range(args.e2e_steps):
            t_s = time.perf_counter()
            e2e_pass("nosync" if os.environ.get("KC_E2E_TRACE") else False)
            e2e_wall.append(round((time.perf_counter() - t_s) * 1e3, 2))  # (the pass ends synchronised: finalize returns counters)
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        # the sink really holds the result: the counts add up to the stage's own statistic
        got = int(res_counts[:final_local].astype(np.uint64).sum())
        assert got == int(stats["sum_counts"]), f"result sink: counts sum {got} != {stats['sum_counts']}"
        st.set_result_sink(None)
        e2e = {"value": occ_all * args.e2e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": int(sum_over_ranks(float(h2d_bytes))),
               "d2h_bytes_per_step": int(sum_over_ranks(float(d2h_bytes))), "steps": args.e2e_steps, "ms_per_step": dt / args.e2e_steps * 1e3, "step_wall_ms": e2e_wall,
               "api": "kc_set_result_sink(host pinned) + kc_add_reads(host pinned) -> kc_parse [-> peer-memory pull or NCCL all-to-all] -> kc_insert -> kc_finalize (survivors land in the sink)"}

    red = None
    if dist is not None:
        from mhm2_kmer_analysis_v2_b200.exchange import reduce_stats

        red = reduce_stats(stats)
    # ---- CPU baseline (rank 0, single-GPU runs only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = min(host_cores(), 128)
        sample = min(n, args.cpu_sample_reads or 100_000 * cores)
        st.close()
        _, _, cpu = cpu_reference_run(k, sample, 1, 0, reads=reads)

    if rank == 0:
        g = red or stats
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic", "config": config_dict(args, world),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "kernel": "insert_kernel (HBM table)" if (args.filter or args.global_table) else "count_kernel (bin-resident insert + purge + extension resolution)",
                         "peak_source": peak_src, "b_alg_bytes_per_kmer": b_alg(k),
                         "kmers_per_launch": inserted, "ms_per_launch": ms_insert_kernel},
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "phases_ms_per_step": {kk: (v / args.steps if kk != "exchange_s" else v / args.steps * 1e3) for kk, v in phase.items()},
            "table": {"unique_kmers": int(g["unique_kmers"]), "final_kmers": int(g["final_kmers"]), "sum_counts": int(g["sum_counts"]),
                      "table_bytes": int(g["table_bytes"]), "supermer_bytes": int(g["supermer_bytes_out"]),
                      "local_bins": int(stats["local_bins"]), "split_passes": int(stats["split_passes"]),
                      "result_bytes": int(stats["result_bytes"])},
            "read_gen_s": t_gen, "lookup": lookup, "step_wall_ms": [round(x, 2) for x in step_wall],
        }
        if world > 1:
            xs = phase["exchange_s"] / args.steps
            if args.exchange == "pull":
                # the transfer happens inside the refine kernel: time = insert phase minus the count kernel
                xs = (phase["insert"] - phase["insert_kernel"]) / args.steps / 1e3
                rec_b = stats["supermer_bytes_out"] / max(stats["supermers_out"], 1)
                # records that crossed NVLink: what this owner consumed minus its own share (owners are balanced by the hash)
                xpulled = stats["supermers_in"] * rec_b * (world - 1) / world * args.steps
                line["exchange"] = {"mode": "pull: refine kernel reads the peers' bins over NVLink (CUDA IPC), no staging",
                                    "bytes_pulled_per_gpu_per_step_approx": xpulled / args.steps, "seconds_refine_incl_local_bins": xs,
                                    "gbps_per_gpu_per_direction": (xpulled / args.steps) / xs / 1e9 if xs > 0 else None,
                                    "nvlink_peak_gbps": 900.0, "owner_imbalance": red.get("owner_imbalance") if red else None}
            else:
                line["exchange"] = {"mode": "nccl all-to-all of owner regions (slack between bins travels too)",
                                    "bytes_sent_per_gpu_per_step": xbytes / args.steps, "seconds": xs,
                                    "gbps_per_gpu_per_direction": (xbytes / args.steps) / xs / 1e9 if xs > 0 else None,
                                    "nvlink_peak_gbps": 900.0, "owner_imbalance": red.get("owner_imbalance") if red else None}
            line["table"]["checksum_sum_over_ranks"] = checksum_all
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
