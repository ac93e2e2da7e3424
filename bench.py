#!/usr/bin/env python
"""Headline benchmark of the k-mer analysis stage (BASELINE.json metric: counted k-mers per second).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload C1..C5 --k K --m M --filter 0|1]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Headline workload = BASELINE config 2 ("C2"): seeded synthetic metagenome mix, 100 random genomes of 1-5 Mbp with
lognormal(0,1) abundance, 20 M x 150 bp reads PER GPU (weak scaling: rank r takes reads [r*20M, (r+1)*20M) of the one seeded
stream), 0.5 % substitutions, k = 33 (two key words), singleton filter off, min depth 2. The other BASELINE configs
(C1 k=21; C2 k=21/55; C3 paired k=21/33/55; C4 k=77/99 with m=15 and m=27; C5 1 % error, filter on / off) are run at reduced
steps after the headline and reported under "configs" in the same JSON line (--configs none|auto|all).

One step = one complete pass of the hot path over the rank's reads: kc_reset, supermer extraction (kc_parse), the owner
exchange when N > 1 (peer-memory pull, copy engines or NCCL all-to-all), refine + count (kc_insert), kc_finalize.
  parity_check  before anything is timed: a small seeded read set goes through the SAME path (exchange mode, table variant,
           k, m) on all ranks; the order-independent result checksum summed over the owners must equal the CPU checker's
           (oracle/_ref/kcount_ref = the reference's own sources, else oracle/libkcount_oracle.so). Mismatch = exit 1.
  value    reads resident in HBM at the start of the step, result left in HBM; timed on the device (CUDA events on the
           stage's stream, host gaps included), max over ranks.
  e2e      same stage through the C ABI with HOST buffers: pinned reads -> kc_add_reads (H2D) -> run -> survivors into a
           pinned host sink (D2H); wall clock between device syncs. --e2e-depth 2 (default): two stage objects, two host
           threads, so the H2D of one job overlaps the D2H of the other (PCIe is full duplex); depth 1 = strictly serial.
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

READ_LEN = 150
METRIC = "counted_kmers_per_sec"
UNIT = "kmers/s"

# BASELINE.json configs. reads = per GPU (weak scaling) unless total_reads is set (C4: "the 20M-read mix at
# 2/4/8 GPUs" is one fixed read set split over the GPUs)
WORKLOADS = {
    "C1": dict(k=21, reads_per_gpu=1_000_000, desc="1 M x 150 bp reads from one random 5 Mbp genome, 0.5% substitutions, seed 1001"),
    "C2": dict(k=33, reads_per_gpu=20_000_000, desc="synthetic metagenome mix, 100 genomes 1-5 Mbp lognormal(0,1), 0.5% substitutions, seed 2002"),
    "C3": dict(k=21, reads_per_gpu=25_000_000, desc="ArcticSynth-shaped paired reads (insert ~N(300,30)), 100-genome mix, 0.5% substitutions, seed 3003"),
    "C4": dict(k=77, reads_per_gpu=20_000_000, total_reads=20_000_000, desc="long-k path on the 20 M-read C2 mix (multi-word keys), seed 2002"),
    "C5": dict(k=21, reads_per_gpu=20_000_000, desc="high-error reads: the C2 community at 1% substitutions, seed 5005"),
}


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


def profile_traffic(k, kernel_variant):
    """DRAM bytes per counted k-mer of the dominant kernel from the committed `ncu --set full` capture of this round
    (profiles/r02_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of one launch / its k-mers), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
            tj = json.load(f)
        e = tj.get(f"{kernel_variant}_k{k}")
        return (e["dram_bytes_per_counted_kmer"], e.get("source")) if e else (None, None)
    except Exception:
        return None, None


def config_dict(spec, n_gpus):
    w = WORKLOADS[spec["workload"]]
    return {
        "workload": f"{spec['workload']}: {w['desc']}, {spec['reads_per_gpu']} x {READ_LEN} bp reads per GPU, k={spec['k']}, kcount stage only",
        "k": spec["k"], "m": spec["m"] or "reference rule clamp(2k/3+1,15,27)", "reads_per_gpu": spec["reads_per_gpu"], "read_len": READ_LEN,
        "n_gpus": n_gpus, "filter": bool(spec["filter"]), "global_table": bool(spec["global_table"]), "dmin_thres": 2,
        "parallelism": f"owner-partitioned x{n_gpus}" if n_gpus > 1 else "single GPU",
        "l2": "inputs (reads, supermer records, result: GBs per GPU) each far larger than the 126 MB L2; no explicit flush",
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


def shm_dir():
    return "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()


def cpu_reference_run(workload, k, sample_reads, n_steps, n_warmup, reads=None, total_reads=None):
    """Times the reference CPU kcount on the first `sample_reads` reads of the workload. Returns (mean seconds per pass,
    counted k-mers, dict describing the run). Uses oracle/_ref/kcount_ref when present (kind "reference"), else the oracle
    restatement (kind "port"). Only libkc_synth.so (the read generator) and oracle/ are touched: no product code."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O

    from mhm2_kmer_analysis_v2_b200 import reads as R

    if reads is None:
        reads = R.generate(workload, sample_reads)
    elif len(reads) > sample_reads:
        reads = reads.slice(0, sample_reads)
    occ = reads.counted_kmers(k)
    times = []
    if os.path.exists(O.REF_BIN):
        cores = min(host_cores(), 128)
        path = os.path.join(shm_dir(), f"kc_bench_reads_{os.getpid()}.bin")
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
    # a prefix of the read stream is shallower than the full set (fewer repeats per distinct k-mer): say how much
    cov = f"; prefix = {100.0 * len(reads) / total_reads:.1f}% of the workload's reads, i.e. that fraction of its coverage -- not like-for-like, a stated CPU baseline" if total_reads else ""
    info = {"value": occ / mean, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"first {len(reads)} reads of the workload ({occ} counted k-mers), {how}; {mean:.2f} s per pass{cov}"}
    return mean, occ, info


def spec_from_args(args, world):
    w = WORKLOADS[args.workload]
    k = args.k or w["k"]
    n = args.reads_per_gpu or (w["total_reads"] // world if "total_reads" in w else w["reads_per_gpu"])
    spec = dict(workload=args.workload, k=k, m=args.m, filter=int(args.filter), global_table=bool(args.global_table), reads_per_gpu=n,
                local_bins=args.local_bins, record_bytes=args.record_bytes, table_slots=args.table_slots)
    if args.n_genomes:
        spec["synth_over"] = {"n_genomes": args.n_genomes}
    return spec


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = min(host_cores(), 128)
    spec = spec_from_args(args, args.gpus)
    total = spec["reads_per_gpu"] * args.gpus
    sample = min(total, args.cpu_sample_reads or 100_000 * cores)
    mean, occ, info = cpu_reference_run(args.workload, spec["k"], sample, args.steps, args.warmup, total_reads=total)
    line = {
        "impl": "reference", "metric": METRIC, "value": info["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": mean * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic", "config": config_dict(spec, args.gpus), "cpu_baseline": info,
        "e2e": {"value": info["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------ GPU arm
class Env:
    """Process-wide state of the GPU arm: ranks, collectives."""

    def __init__(self, args):
        import torch

        self.torch = torch
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus:
            if self.world == 1 and args.gpus > 1:
                raise SystemExit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks (one process per GPU)")
            raise SystemExit(f"WORLD_SIZE {self.world} != --gpus {args.gpus}")
        torch.cuda.set_device(self.local_rank)
        self.dist = None
        self.groups = [None, None]
        if self.world > 1:
            import datetime

            import torch.distributed as dist

            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank), timeout=datetime.timedelta(seconds=600))
            self.dist = dist
            # control-plane groups of the pipelined end-to-end arm: each host thread drives its own job with its own
            # (CPU) group, so their small all-gathers never interleave on one communicator
            self.groups = [dist.new_group(backend="gloo", timeout=datetime.timedelta(seconds=600)) for _ in range(2)]

    def barrier(self, *stages):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()
        for st in stages:
            st.sync()

    def max(self, x):
        if self.dist is None:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def min(self, x):
        return -self.max(-x)

    def sum(self, x):
        if self.dist is None:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t)
        return float(t.item())

    def sum_u64(self, v):
        """sum modulo 2^64 over the ranks (order-independent checksums)"""
        if self.dist is None:
            return int(v)
        parts = [None] * self.world
        self.dist.all_gather_object(parts, int(v))
        return sum(parts) % (1 << 64)


def make_stage(K, env, spec, est_kmers, **kw):
    return K.KcountStage(spec["k"], m=spec["m"], rank=env.rank, n_ranks=env.world, device=env.local_rank, est_kmers=est_kmers,
                         use_filter=bool(spec["filter"]), local_bins=kw.pop("local_bins", spec["local_bins"]), record_bytes=spec["record_bytes"],
                         table_slots=spec["table_slots"], global_table=spec["global_table"], **kw)


def one_pass(env, st, exchange, group=None):
    """reset -> parse -> exchange -> insert -> finalize (-> release): one step of the hot path on this rank."""
    from mhm2_kmer_analysis_v2_b200.exchange import exchange_supermers, exchange_supermers_pull, release_peers

    st.reset()
    st.parse()
    xb = (0, 0, 0.0)
    if env.world > 1:
        xb = exchange_supermers(st, group) if exchange == "nccl" else exchange_supermers_pull(st, group)
    st.insert()
    fin = st.finalize()
    if env.world > 1 and exchange != "nccl":
        release_peers(st, group)  # nobody parses again while a peer may still be reading this rank's bins
    return fin, xb


def parity_reads_per_rank(args, world):
    return max(args.parity_reads // world, 1000)


def parity_reference(args, spec, world):
    """The CPU checker's answer for the parity pass: (checksum of D1, survivors, sum of counts, counted k-mers). Runs on rank 0
    BEFORE torch / CUDA are initialised: it forks (oracle/_ref/kcount_ref, one rank per host core) and churns host memory,
    neither of which should happen next to a live CUDA context that is about to be timed."""
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O

    from mhm2_kmer_analysis_v2_b200 import reads as R

    k = spec["k"]
    total = parity_reads_per_rank(args, world) * world
    allr = R.generate(spec["workload"], total, **spec.get("synth_over", {}))
    t0 = time.perf_counter()
    if os.path.exists(O.REF_BIN):
        path = os.path.join(shm_dir(), f"kc_parity_{os.getpid()}")
        allr.save(path + ".reads")
        try:
            s = O.run_reference(path + ".reads", k, path, ranks=min(host_cores(), 32))
            d1 = O.load_dump(path + ".d1")
            want = (O.result_checksum(d1, False), len(d1), int(d1.counts.astype(np.uint64).sum()), int(s["occurrences"]))
        finally:
            for suf in (".reads", ".d1"):
                if os.path.exists(path + suf):
                    os.unlink(path + suf)
        checker = "oracle/_ref/kcount_ref (reference sources)"
    else:
        o = O.Oracle(k, slots_per_rank=max(4 * allr.counted_kmers(k), 1 << 16))
        o.add_reads(allr.bytes, allr.offsets, allr.lens)
        o.finalize()
        fin = o.result(False)
        want = (o.checksum(False), len(fin), int(fin.counts.astype(np.uint64).sum()), o.counted_kmers())
        o.close()
        checker = "oracle/libkcount_oracle.so (C restatement)"
    return {"want": want, "checker": checker, "checker_s": round(time.perf_counter() - t0, 2)}


def parity_check(K, env, args, spec, ref):
    """A small seeded read set through the same path on every rank, against the CPU checker's answer (rank 0 holds it)."""
    import numpy as np

    from mhm2_kmer_analysis_v2_b200 import reads as R

    k = spec["k"]
    n_rank = parity_reads_per_rank(args, env.world)
    total = n_rank * env.world
    mine = R.generate(spec["workload"], n_rank, first=env.rank * n_rank, **spec.get("synth_over", {}))
    # enough fine bins that the two-level partition (parse -> refine -> count, sliced) is the path under test, as at full size
    bins = spec["local_bins"] or (0 if (spec["filter"] or spec["global_table"]) else 131072)
    st = make_stage(K, env, spec, est_kmers=int(np.maximum(mine.lens.astype(np.int64) - k + 1, 0).sum()), local_bins=bins)
    if env.world > 1 and args.exchange != "nccl":
        st.set_exchange_mode(args.exchange)
    st.add_reads(mine)
    for _ in range(2):  # twice: buffers, IPC mappings and layouts are reused from the second pass on
        stats, _ = one_pass(env, st, args.exchange)
    cs = env.sum_u64(st.checksum(False))
    final = int(env.sum(float(stats["final_kmers"])))
    sum_counts = int(env.sum(float(stats["sum_counts"])))
    inserted = int(env.sum(float(stats["inserted_kmers"])))
    fine_bins = int(stats["local_bins"])
    st.close()
    del mine
    out = None
    if env.rank == 0:
        got, want = (cs, final, sum_counts, inserted), tuple(ref["want"])
        mode = "single" if env.world == 1 else args.exchange
        out = {"reads": total, "k": k, "mode": mode, "fine_bins": fine_bins, "survivors": final, "checksum": f"{cs:016x}",
               "checker": ref["checker"], "checker_s": ref["checker_s"], "ok": got == want}
        if got != want:
            out["got"], out["want"] = list(got), list(want)
    if env.dist is not None:
        box = [out]
        env.dist.broadcast_object_list(box, src=0)
        out = box[0]
    return out


def load_reads(K, env, spec, first=None):
    """This rank's reads of the workload, generated straight into pinned host memory."""
    import numpy as np

    from mhm2_kmer_analysis_v2_b200.reads import PinnedBuffer

    n = spec["reads_per_gpu"]
    pin_reads = PinnedBuffer(n * READ_LEN)
    syn = K.Synth(**dict(K.CONFIGS[spec["workload"]]["synth"], **spec.get("synth_over", {})))
    t0 = time.perf_counter()
    reads = syn.generate(env.rank * n if first is None else first, n, out_bytes=pin_reads.array)
    syn.close()
    pin_off, pin_len = PinnedBuffer(n * 8), PinnedBuffer(n * 2)
    pin_off.view(np.uint64, n)[:] = reads.offsets
    pin_len.view(np.uint16, n)[:] = reads.lens
    nb = len(reads.bytes)
    reads = K.ReadSet.__new__(K.ReadSet)
    reads.bytes, reads.offsets, reads.lens = pin_reads.array[:nb], pin_off.view(np.uint64, n), pin_len.view(np.uint16, n)
    reads._pins = (pin_reads, pin_off, pin_len)
    return reads, time.perf_counter() - t0


def free_reads(reads):
    pins = reads._pins
    reads.bytes = reads.offsets = reads.lens = None
    reads._pins = None
    for p in pins:
        p.free()


def measure(K, env, args, spec, steps, warmup, headline):
    """Device-resident arm of one configuration (+ lookups, end to end and the CPU baseline for the headline)."""
    import numpy as np

    torch = env.torch
    k, n = spec["k"], spec["reads_per_gpu"]
    reads, t_gen = load_reads(K, env, spec)
    occ_local = reads.counted_kmers(k)
    raw_local = int(np.maximum(reads.lens.astype(np.int64) - k + 1, 0).sum())
    h2d_bytes = len(reads.bytes) + n * 10
    st = make_stage(K, env, spec, est_kmers=raw_local)
    if env.world > 1 and args.exchange != "nccl":
        st.set_exchange_mode(args.exchange)
    st.add_reads(reads)
    st.sync()
    sampler = None
    if headline and env.rank == 0:
        # the clock sampler (an nvidia-smi process) starts before the warm-up: its start-up perturbs the driver for a moment,
        # which must not fall into the timed region; it then samples every 200 ms through warm-up and timed steps
        sampler = ClockSampler(env.local_rank)
        sampler.start()
        sampler.wait_first_sample()
    env.barrier(st)
    for _ in range(warmup):
        stats, _ = one_pass(env, st, args.exchange)
    checksum_warm = st.checksum(False) if warmup > 0 else None
    launches0 = st.stats()["kernel_launches"]
    names = ("parse", "parse_kernel", "insert", "insert_kernel", "refine_kernel", "table_clear", "finalize", "exchange_copy")
    phase = {nm: 0.0 for nm in names}
    phase["exchange_s"] = 0.0
    xbytes = 0
    invariants = set()
    env.barrier(st)
    st.timer_mark(0)
    step_wall = []
    for _ in range(steps):
        t_s = time.perf_counter()
        stats, xb = one_pass(env, st, args.exchange)
        step_wall.append((time.perf_counter() - t_s) * 1e3)  # (finalize returns the counters: every step ends synchronised)
        for nm in names:
            phase[nm] += stats["ms_" + nm]
        phase["exchange_s"] += xb[2]
        xbytes += xb[0]
        # (with the filter on, unique_kmers = keys ADMITTED to the table: a tag collision admits a singleton, which is then
        # counted exactly and purged -- how many do depends on the order of arrival, the result does not)
        invariants.add((stats["final_kmers"], stats["sum_counts"], 0 if spec["filter"] else stats["unique_kmers"], stats["ge2_kmers"],
                        stats["max_count"], stats["inserted_kmers"]))
    st.timer_mark(1)
    env.barrier(st)
    ms_total = env.max(st.timer_elapsed_ms(0, 1))
    clocks = sampler.stop() if sampler else None
    launches = st.stats()["kernel_launches"] - launches0
    checksum = st.checksum(False)
    if checksum_warm is not None and checksum != checksum_warm:
        raise SystemExit(f"bench: the result changed between steps (checksum {checksum_warm:016x} -> {checksum:016x})")
    stable = {"steps_compared": steps, "identical": len(invariants) == 1, "what": "final, sum_counts, unique, ge2, max_count, inserted of every timed step; "
              "result checksum after the warm-up and after the last timed step"}
    if len(invariants) != 1:
        stable["sets_seen"] = sorted(invariants)
        print(f"bench: WARNING: counters differ between steps: {sorted(invariants)}", file=sys.stderr)
    occ_all = env.sum(float(occ_local))
    value = occ_all * steps / (ms_total / 1e3)
    ms_insert_kernel = phase["insert_kernel"] / steps
    inserted = stats["inserted_kmers"]  # k-mers this rank's insert launches processed
    if env.world == 1 and inserted != occ_local:
        raise SystemExit(f"bench: {inserted} k-mers inserted, {occ_local} expected")
    peak, peak_src = load_peaks()
    achieved = b_alg(k) * inserted / (ms_insert_kernel / 1e3) / 1e9 if ms_insert_kernel > 0 else 0.0
    hbm_table = bool(spec["filter"] or spec["global_table"])
    per_kmer, traffic_src = profile_traffic(k, "insert" if hbm_table else "count")
    final_local = stats["final_kmers"]
    free_b, total_b = torch.cuda.mem_get_info()
    rec = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": env.world, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_total / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic", "config": config_dict(spec, env.world),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": per_kmer * inserted if per_kmer else None, "traffic_source": traffic_src,
                     "kernel": "insert_kernel (HBM table)" if hbm_table else "count_kernel (bin-resident insert + purge + extension resolution)",
                     "peak_source": peak_src, "b_alg_bytes_per_kmer": b_alg(k), "kmers_per_launch": inserted, "ms_per_launch": ms_insert_kernel,
                     "launches_per_step": int(stats["pipeline_slices"]) if not hbm_table else int(stats["insert_kernel_launches"])},
        "gpu_launches": int(launches), "clocks": clocks,
        "phases_ms_per_step": {kk: (v / steps if kk != "exchange_s" else v / steps * 1e3) for kk, v in phase.items()},
        "checksum": f"{env.sum_u64(checksum):016x}", "result_stability": stable,
        "hbm_bytes_in_use": int(total_b - free_b),
        "read_gen_s": t_gen, "step_wall_ms": [round(x, 2) for x in step_wall],
    }
    red = None
    if env.dist is not None:
        from mhm2_kmer_analysis_v2_b200.exchange import reduce_stats

        red = reduce_stats(stats)
    g = red or stats
    rec["table"] = {"unique_kmers": int(g["unique_kmers"]), "final_kmers": int(g["final_kmers"]), "sum_counts": int(g["sum_counts"]),
                    "table_bytes": int(g["table_bytes"]), "filter_bytes": int(stats["filter_bytes"]), "supermer_bytes": int(g["supermer_bytes_out"]),
                    "local_bins": int(stats["local_bins"]), "split_passes": int(stats["split_passes"]), "result_bytes": int(stats["result_bytes"]),
                    "insert_passes": int(stats["insert_kernel_launches"]) if hbm_table else 1}
    if env.world > 1:
        rec_b = stats["supermer_bytes_out"] / max(stats["supermers_out"], 1)
        # records that crossed NVLink: what this owner consumed minus its own share (owners are balanced by the hash)
        pulled = stats["supermers_in"] * rec_b * (env.world - 1) / env.world
        if args.exchange == "pull":
            xs = phase["refine_kernel"] / steps / 1e3  # the transfer happens inside the refine kernel
            how = "pull: refine kernel reads the peers' bins over NVLink (CUDA IPC) slice by slice beside the count kernel, no staging"
        elif args.exchange == "copy":
            xs = phase["exchange_copy"] / steps / 1e3
            pulled = stats["exchange_copy_bytes"]
            how = "copy: copy engines move the peers' regions (CUDA IPC) slice by slice beside the refine / count kernels"
        else:
            xs = phase["exchange_s"] / steps
            pulled = xbytes / steps
            how = "nccl all-to-all of owner regions (slack between bins travels too), then refine + count"
        rec["exchange"] = {"mode": how, "bytes_per_gpu_per_step": pulled, "seconds": xs,
                           "gbps_per_gpu_per_direction": pulled / xs / 1e9 if xs > 0 else None, "nvlink_peak_gbps": 900.0,
                           "nvlink_frac": pulled / xs / 1e9 / 900.0 if xs > 0 else None,
                           "owner_imbalance": red.get("owner_imbalance") if red else None}
    rec["reference_pin"] = check_full_pin(spec, env, g, rec["checksum"])
    if not headline:
        st.close()
        free_reads(reads)
        return rec

    # ---- the transport north_star names literally, beside the default: same stage, NCCL all-to-all, 2 steps
    if env.world > 1 and args.exchange != "nccl":
        for _ in range(2):
            one_pass(env, st, "nccl")
        env.barrier(st)
        st.timer_mark(2)
        xs = 0.0
        for _ in range(2):
            s2, xb = one_pass(env, st, "nccl")
            xs += xb[2]
        st.timer_mark(3)
        env.barrier(st)
        ms2 = env.max(st.timer_elapsed_ms(2, 3)) / 2
        rec["exchange_nccl"] = {"ms_per_step": ms2, "value": occ_all / (ms2 / 1e3), "all_to_all_seconds": xs / 2, "bytes_sent_per_gpu_per_step": xb[0],
                                "gbps_per_gpu_per_direction": xb[0] / (xs / 2) / 1e9 if xs > 0 else None,
                                "checksum": f"{env.sum_u64(st.checksum(False)):016x}"}
        if rec["exchange_nccl"]["checksum"] != rec["checksum"]:
            raise SystemExit("bench: the NCCL exchange produced another result than the default exchange")
        one_pass(env, st, args.exchange)  # (the result the lookups below query)

    # ---- batched lookups against the device-resident result (what the next stage asks the table)
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
        rec["lookup"] = {"queries": nq, "queries_per_sec": nq / t_q, "ms_per_batch": t_q * 1e3, "index_build_ms": max(t_first - t_q, 0.0) * 1e3,
                         "entries": int(final_local), "hit_fraction_asked": 0.5, "miss_found": int((lc[1::2] != 0).sum().item()),
                         "where": "queries and answers resident in HBM (kc_lookup with device pointers)"}
        del q, idx, lc, ll, lr, rd

    # ---- end-to-end arm: host reads in, host result out, through the C ABI
    if not args.no_e2e:
        try:
            rec["e2e"] = e2e_arm(K, env, args, spec, st, reads, stats, occ_all, h2d_bytes, raw_local)
        except Exception as e:  # noqa: BLE001 -- the device-resident headline stands; the failure is reported, and all ranks agree on it
            rec["e2e"] = {"error": str(e)[:300]}
            print(f"bench: end-to-end arm failed: {e}", file=sys.stderr)
    # ---- CPU baseline (rank 0, single-GPU runs only)
    st.close()
    if env.rank == 0 and env.world == 1 and not args.no_cpu_baseline:
        cores = min(host_cores(), 128)
        sample = min(n, args.cpu_sample_reads or 100_000 * cores)
        _, _, rec["cpu_baseline"] = cpu_reference_run(spec["workload"], k, sample, 1, 0, reads=reads, total_reads=n)
    else:
        rec["cpu_baseline"] = None
    free_reads(reads)
    return rec


def check_full_pin(spec, env, g, checksum):
    """The reference's own result for this exact configuration, when tests/golden/full_config_pins.json holds one
    (tools/pin_full_config.py ran oracle/_ref/kcount_ref on all reads of the workload)."""
    try:
        with open(os.path.join(ROOT, "tests", "golden", "full_config_pins.json")) as f:
            pins = json.load(f)["pins"]
    except Exception:
        return None
    for p in pins:
        if (p["workload"], p["k"], p["n_reads"]) == (spec["workload"], spec["k"], spec["reads_per_gpu"] * env.world) and not spec.get("synth_over"):
            got = {"entries": int(g["unique_kmers"]), "survivors": int(g["final_kmers"]), "sum_count": int(g["sum_counts"]), "checksum_d1": checksum}
            want = {kk: p[kk] for kk in got}
            if spec["filter"]:
                got.pop("entries"), want.pop("entries")
            ok = got == want
            if not ok:
                print(f"bench: result differs from the reference pin: {got} != {want}", file=sys.stderr)
            return {"ok": ok, "source": "tests/golden/full_config_pins.json: oracle/_ref/kcount_ref (reference sources) on all reads of this configuration",
                    "reads": p["n_reads"], "survivors": p["survivors"]}
    return None


def e2e_arm(K, env, args, spec, st0, reads, stats, occ_all, h2d_bytes, raw_local):
    import numpy as np

    from mhm2_kmer_analysis_v2_b200.reads import PinnedBuffer

    depth = max(1, min(args.e2e_depth, 2))
    if env.world > 1 and args.exchange == "nccl":
        depth = 1  # the data all-to-all runs on the one NCCL communicator: one job at a time
    if depth == 2:
        # a second stage object needs what the first one holds (bins, fine bins, receive buffer, result): only with room for it
        free_b, total_b = env.torch.cuda.mem_get_info()
        if env.min(float(free_b) - 1.3 * float(total_b - free_b)) < 0:
            depth = 1
    groups = env.groups if not (env.world > 1 and args.exchange == "nccl") else [None, None]
    final_local = stats["final_kmers"]
    N = st0.N
    d2h_bytes = final_local * (8 * N + 4)
    stages = [st0]
    if depth == 2:
        st1 = make_stage(K, env, spec, est_kmers=raw_local)
        if env.world > 1 and args.exchange != "nccl":
            st1.set_exchange_mode(args.exchange)
        stages.append(st1)
    sinks = []
    for st in stages:
        pin = PinnedBuffer(max(d2h_bytes, 1) + 64)
        keys = pin.view(np.uint64, final_local * N)
        counts = pin.view(np.uint16, final_local, offset=final_local * N * 8)
        left = pin.view(np.uint8, final_local, offset=final_local * (N * 8 + 2))
        right = pin.view(np.uint8, final_local, offset=final_local * (N * 8 + 3))
        st.set_result_sink(keys, counts, left, right)
        sinks.append((pin, keys, counts, left, right))
    errors = []

    def job(t, n_pass, walls):
        try:
            torch = env.torch
            torch.cuda.set_device(env.local_rank)
            st = stages[t]
            for _ in range(n_pass):
                t_s = time.perf_counter()
                st.clear_reads()
                st.add_reads(reads)  # H2D from pinned memory, asynchronous on the stage's input stream
                s, _ = one_pass(env, st, args.exchange, groups[t])  # the survivors stream into the sink while the count kernel runs
                if s["final_kmers"] != final_local:
                    raise RuntimeError("result size changed between passes")
                walls.append(round((time.perf_counter() - t_s) * 1e3, 2))
        except BaseException as e:  # noqa: BLE001 -- re-raised on the main thread
            errors.append(e)

    def run(n_each):
        walls = [[] for _ in stages]
        ths = [threading.Thread(target=job, args=(t, n_each, walls[t])) for t in range(len(stages))]
        t0 = time.perf_counter()
        for th in ths:
            th.start()
        for th in ths:
            th.join()
        if errors:
            raise errors[0]
        return time.perf_counter() - t0, walls

    run(2)  # warm-up: buffer pools and bin layouts settle over the first two passes of every stage
    n_each = max(args.e2e_steps // depth, 1)
    env.barrier(*stages)
    dt, walls = run(n_each)
    env.barrier(*stages)
    dt = env.max(dt)
    n_jobs = n_each * depth
    # the sinks really hold the result: the counts add up to the stage's own statistic
    for (_, _, counts, _, _) in sinks:
        got = int(counts[:final_local].astype(np.uint64).sum())
        assert got == int(stats["sum_counts"]), f"result sink: counts sum {got} != {stats['sum_counts']}"
    # ---- the same pass starting from the reference's real in-memory reads: 16-byte PackedRead headers
    # pointing into 4 MB allocator blocks -> kc_add_packed_reads (host threads gather into pinned staging, piece by piece)
    packed = None
    if env.world == 1 and not args.no_packed:
        layout = K.PackedReadsLayout(reads)
        st = stages[0]
        t_add, t_all = [], []
        for it in range(3):
            t_s = time.perf_counter()
            st.clear_reads()
            st.add_packed_reads(layout)
            t_a = time.perf_counter()
            s, _ = one_pass(env, st, args.exchange, groups[0])
            t_e = time.perf_counter()
            if s["final_kmers"] != final_local or s["sum_counts"] != stats["sum_counts"]:
                raise RuntimeError("kc_add_packed_reads: another result than kc_add_reads")
            if it:
                t_add.append((t_a - t_s) * 1e3), t_all.append((t_e - t_s) * 1e3)
        packed = {"ms_per_step": sum(t_all) / len(t_all), "ms_add_packed_reads": sum(t_add) / len(t_add), "allocator_blocks": layout.n_blocks,
                  "value": occ_all / (sum(t_all) / len(t_all) / 1e3), "jobs_in_flight": 1,
                  "what": "kc_add_packed_reads(vector<PackedRead> headers -> 4 MB blocks, pageable host memory) -> parse -> insert -> finalize into the pinned "
                          "sink; the gather of the bytes into pinned staging by host threads is inside the time"}
        del layout
    for st in stages:
        st.set_result_sink(None)
    for st in stages[1:]:
        st.close()
    h2d_all, d2h_all = env.sum(float(h2d_bytes)), env.sum(float(d2h_bytes))
    ms = dt / n_jobs * 1e3
    out = {"value": occ_all * n_jobs / dt, "unit": UNIT, "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(d2h_all),
           "steps": n_jobs, "ms_per_step": ms, "jobs_in_flight": depth, "job_wall_ms": walls,
           "pcie_gbps_per_gpu": {"h2d": h2d_bytes / (ms / 1e3) / 1e9, "d2h": d2h_bytes / (ms / 1e3) / 1e9},
           "from_packed_reads_container": packed,
           "api": "kc_set_result_sink(host pinned) + kc_add_reads(host pinned) -> kc_parse [-> exchange] -> kc_insert -> kc_finalize (survivors land in the "
                  "sink)" + ("; two stage objects driven by two host threads, so one job's input copy overlaps the other's result copy" if depth == 2 else "")}
    for s in sinks:
        s[0].free()
    return out


def merge_cpu_baseline(ps, st=None, sample_pairs=200_000):
    """The same pair loop on one host core: oracle/merge_oracle.c (plain-C restatement of src/merge_reads.cpp:466-650; the
    reference runs this loop on one rank per core) on a prefix of the pairs -- the cpu_baseline leg, the one place besides the
    parity pass where bench.py may execute oracle/. Also a checker: its merged / ambiguous counts on the full set are not
    compared here (tests/test_merge.py does that), only reported for the sample."""
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O

    n = min(sample_pairs, len(ps))
    end = int(ps.offsets[2 * n - 1]) + int(ps.lens[2 * n - 1]) if n else 0
    t0 = time.perf_counter()
    _, _, _, ctr = O.merge_pairs(ps.seq[:end], ps.qual[:end], ps.offsets[: 2 * n], ps.lens[: 2 * n])
    dt = time.perf_counter() - t0
    return {"value": n / dt if dt > 0 else None, "unit": "pairs/s", "cores": 1, "kind": "port", "sample": f"first {n} pairs, oracle/merge_oracle.c",
            "merged_in_sample": ctr["merged"]}


def merge_bench(K, env, n_pairs=2_000_000):
    """Read-pair merging in front of the stage (kc_merge.cu replaces the pair loop of merge_reads(),
    src/merge_reads.cpp:466-650): seeded synthetic pairs (C3's community, 150 bp mates, insert N(220, 40) so that most mates
    overlap) as FASTQ characters in host memory -> kc_merge_pairs -> PackedRead bytes in host memory. Parity of this path:
    tests/test_merge.py (reference goldens). Reported beside the headline, not part of it."""
    import numpy as np

    rs = K.generate("C3", 2 * n_pairs, n_genomes=20, insert_mean=220.0, insert_sd=40.0)
    lut_b = np.frombuffer(b"ACGTNNNN", np.uint8)
    seq = lut_b[rs.bytes & 7]
    qual = (rs.bytes >> 3) + np.uint8(33)
    ps = K.PairSet(seq, qual, rs.offsets, rs.lens)
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        reads, st = K.merge_pairs(ps, device=env.local_rank if hasattr(env, "local_rank") else 0)
        wall = time.perf_counter() - t0
        if best is None or st["ms_kernel"] < best[0]["ms_kernel"]:
            best = (st, wall)
    st, wall = best
    ms = st["ms_kernel"]
    cpu = merge_cpu_baseline(ps, st)
    return {"cpu_baseline": cpu, "pairs": n_pairs, "merged": st["merged"], "ambiguous": st["ambiguous"], "kernel_ms": ms, "pairs_per_sec": n_pairs / (ms / 1e3),
            "algorithmic_gbps": 3 * len(seq) / (ms / 1e3) / 1e9, "call_wall_ms": wall * 1e3, "gpu_launches": st["launches"],
            "what": "2 bytes read + 1 written per base; kernel time by CUDA events inside kc_merge_pairs; the call's wall time includes its own "
                    "cudaMalloc / H2D / D2H of pageable host arrays"}


def sub_config_specs(args, world):
    """(label, spec overrides) of the other BASELINE configs this run can hold."""
    base = dict(m=0, filter=0, global_table=False, local_bins=0, record_bytes=0, table_slots=0)
    out = []

    def add(label, workload, k, **kw):
        w = WORKLOADS[workload]
        n = kw.pop("reads_per_gpu", None) or (w["total_reads"] // world if "total_reads" in w else w["reads_per_gpu"])
        out.append((label, dict(base, workload=workload, k=k, reads_per_gpu=n, **kw)))

    if world == 1:
        add("C1_k21", "C1", 21)
        add("C2_k21", "C2", 21)
        add("C2_k55", "C2", 55)
    if world in (1, 2, 4, 8):
        for k in (77, 99):
            add(f"C4_k{k}_m15", "C4", k, m=15)
            add(f"C4_k{k}_m27", "C4", k, m=27)
    if world == 1:
        add("C5_k21_filter_off_resident", "C5", 21)
        add("C5_k21_filter_off_hbm_table", "C5", 21, global_table=True)
        add("C5_k21_filter_on", "C5", 21, filter=1)
    if world == 8 or args.configs == "all":
        for k in (21, 33, 55):
            add(f"C3_k{k}", "C3", k)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS))
    ap.add_argument("--k", type=int, default=0, help="0 = the workload's k")
    ap.add_argument("--m", type=int, default=0, help="minimizer length; 0 = the reference rule clamp(2k/3+1, 15, 27)")
    ap.add_argument("--reads-per-gpu", type=int, default=0, help="0 = the workload's size")
    ap.add_argument("--filter", type=int, default=0, help="1 = two-choice singleton filter on")
    ap.add_argument("--local-bins", type=int, default=0)
    ap.add_argument("--record-bytes", type=int, default=0)
    ap.add_argument("--table-slots", type=int, default=0, help="exact table size in slots (0 = the reference's sizing rule)")
    ap.add_argument("--global-table", action="store_true", help="count in one HBM table (kc_insert.cu) instead of bin-resident tables")
    ap.add_argument("--e2e-steps", type=int, default=4)
    ap.add_argument("--e2e-depth", type=int, default=2, help="jobs in flight in the end-to-end arm (1 = strictly serial)")
    ap.add_argument("--cpu-sample-reads", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--n-genomes", type=int, default=0, help="profiling only: fewer genomes keep the coverage of a smaller read set like the full workload's")
    ap.add_argument("--no-packed", action="store_true", help="skip the end-to-end pass that starts from the reference's PackedReads memory layout")
    ap.add_argument("--no-merge", action="store_true", help="skip the read-pair merging measurement (kc_merge_pairs, beside the headline)")
    ap.add_argument("--no-lookup", action="store_true", help="skip the batched-lookup measurement against the finished result")
    ap.add_argument("--no-parity", action="store_true", help="skip the untimed parity pass against the CPU checker")
    ap.add_argument("--parity-reads", type=int, default=400_000, help="reads (all ranks together) of the parity pass")
    ap.add_argument("--configs", default="auto", choices=["none", "auto", "all"], help="also measure the other BASELINE configs (reduced steps)")
    ap.add_argument("--exchange", default="pull", choices=["pull", "copy", "nccl"],
                    help="multi-GPU: 'pull' = peers' bins mapped with CUDA IPC, the refine kernel reads them over NVLink (>= 4 GPUs: on SMs "
                         "of its own beside the count kernel); 'copy' = copy engines move them slice by slice; 'nccl' = all-to-all of "
                         "the owner regions, then refine")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        print(f"note: warmup {args.warmup} < 3 (timing rules ask for >= 3)", file=sys.stderr)

    if args.impl == "reference":
        return run_reference_arm(args)

    world0 = int(os.environ.get("WORLD_SIZE", "1"))
    ref = None
    if not args.no_parity and int(os.environ.get("RANK", "0")) == 0:
        ref = parity_reference(args, spec_from_args(args, world0), world0)  # before any CUDA context exists in this process

    import mhm2_kmer_analysis_v2_b200 as K

    if K._lib.lib().kc_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the stage has no CPU path (use --impl reference for the CPU arm)")
    env = Env(args)
    spec = spec_from_args(args, env.world)

    parity = None
    if not args.no_parity:
        parity = parity_check(K, env, args, spec, ref)
        if not parity["ok"]:
            if env.rank == 0:
                print(json.dumps({"parity_check": parity}), flush=True)
            raise SystemExit("bench: parity check against the CPU checker FAILED")

    line = measure(K, env, args, spec, args.steps, args.warmup, headline=True)
    line["parity_check"] = parity
    if args.configs != "none":
        cfgs = {}
        for label, sp in sub_config_specs(args, env.world):
            try:
                r = measure(K, env, args, sp, 2, 2, headline=False)
                cfgs[label] = {"value": r["value"], "ms_per_step": r["ms_per_step"], "k": sp["k"], "m": sp["m"] or None, "filter": bool(sp["filter"]),
                               "global_table": bool(sp["global_table"]), "reads_per_gpu": sp["reads_per_gpu"], "steps": 2, "warmup": 2,
                               "phases_ms_per_step": r["phases_ms_per_step"], "roofline_frac": r["roofline"]["frac"],
                               "roofline_kernel": r["roofline"]["kernel"], "b_alg_bytes_per_kmer": r["roofline"]["b_alg_bytes_per_kmer"],
                               "split_passes": r["table"]["split_passes"], "local_bins": r["table"]["local_bins"],
                               "unique_kmers": r["table"]["unique_kmers"], "final_kmers": r["table"]["final_kmers"],
                               "table_bytes": r["table"]["table_bytes"], "filter_bytes": r["table"]["filter_bytes"],
                               "insert_passes": r["table"]["insert_passes"], "hbm_bytes_in_use": r["hbm_bytes_in_use"], "checksum": r["checksum"],
                               "reference_pin_ok": r["reference_pin"]["ok"] if r.get("reference_pin") else None,
                               "exchange": r.get("exchange")}
            except SystemExit:
                raise
            except Exception as e:  # noqa: BLE001 -- a sub-config must not take the headline down; the failure is reported
                cfgs[label] = {"error": str(e)[:300]}
        line["configs"] = cfgs
    if not args.no_merge and env.rank == 0:
        try:
            line["merge_pairs"] = merge_bench(K, env)
        except Exception as e:  # noqa: BLE001 -- reported, not fatal: the step in front of the stage is not the headline
            line["merge_pairs"] = {"error": str(e)[:300]}
    if env.rank == 0:
        print(json.dumps(line), flush=True)
    if env.dist is not None:
        env.dist.barrier()
        env.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
