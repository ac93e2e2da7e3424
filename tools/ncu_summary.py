#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): per-kernel key metrics and the top stalled SASS lines.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [kernel-substring] [n_top]"""
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
want_kernel = sys.argv[2] if len(sys.argv) > 2 else ""
n_top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_red.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_atom.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_atom.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__thread_inst_executed_per_inst_executed.ratio"]
for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    if want_kernel and want_kernel not in name:
        continue
    print("=====", name, "id", r[idx["ID"]])
    for k in KEYS:
        if k in idx:
            print(f"  {k:72s} {r[idx[k]]:>20s} {units[idx[k]]}")
if want_kernel:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{want_kernel}"], capture_output=True, text=True).stdout
    # several kernels may be concatenated: split on header lines
    blocks, cur = [], []
    for line in csv.reader(io.StringIO(src)):
        if line and line[0] == "Kernel Name":
            if cur:
                blocks.append(cur)
            cur = [line]
        else:
            cur.append(line)
    if cur:
        blocks.append(cur)
    for blk in blocks[-1:]:
        print("##", blk[0][1])
        h = blk[1]
        ix = {x: i for i, x in enumerate(h)}
        data = [r for r in blk[2:] if len(r) == len(h)]
        tot = sum(int(r[ix["# Samples"]]) for r in data) or 1
        toti = sum(int(r[ix["Instructions Executed"]]) for r in data)
        print(f"total samples {tot}, warp instructions {toti/1e9:.2f} G")
        stalls = [x for x in h if x.startswith("stall_") and "Not Issued" not in x]
        agg = {x: sum(int(r[ix[x]]) for r in data) for x in stalls}
        print("stall mix:", ", ".join(f"{k[6:]}={v/tot*100:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
        for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]]))[:n_top]:
            s = int(r[ix["# Samples"]])
            st = sorted(((int(r[ix[x]]), x) for x in stalls), reverse=True)[:2]
            print(f"{s/tot*100:5.1f}% inst={int(r[ix['Instructions Executed']])/1e6:9.1f}M thr={r[ix['Avg. Threads Executed']]:>4s} "
                  f"{r[ix['Source']].strip()[:64]:64s} {st[0][1][6:]}:{st[0][0]} {st[1][1][6:]}:{st[1][0]}")
