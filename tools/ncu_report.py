#!/usr/bin/env python
"""One-page summary of an `ncu --set full` report (runs `ncu -i` itself, so it works on the GPU box too).
usage: tools/ncu_report.py rep.ncu-rep [units_per_launch] > profiles/x.txt
units_per_launch: env-us one launch simulates (n_envs * n_us), for the per-env-us figures."""
import csv, subprocess, sys
rep = sys.argv[1]
units = float(sys.argv[2]) if len(sys.argv) > 2 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, unit = rows[0], rows[1]
KEYS = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__block_size", "launch__grid_size",
        "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum",
        "smsp__inst_executed_op_global_st.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores", "sm__cycles_elapsed.avg.per_second"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    u = dict(zip(hdr, unit))
    for k in KEYS:
        if k in d:
            print(f"{k} [{u[k]}]: {d[k]}")
    if units:
        try:
            print(f"warp-instructions per env-us: {float(d['smsp__inst_executed.sum']) / units:.2f}")
            t = float(d["gpu__time_duration.sum"])
            t_s = t * {"us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}.get(u["gpu__time_duration.sum"], 1e-6)
            print(f"env-us per second in this (serialised, cold-cache) launch: {units / t_s:.3e}")
        except Exception:
            pass
    for h in hdr:
        if "issue_stalled" in h and "per_issue_active" in h:
            try:
                if float(d[h]) > 0.05:
                    print("stall per issue:", h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), d[h])
            except ValueError:
                pass
    print()
