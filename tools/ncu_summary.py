"""Print the handful of ncu raw-page metrics the design discussion uses. usage: ncu_summary.py raw.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__waves_per_multiprocessor',
        'launch__occupancy_limit_registers', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores', 'sm__cycles_elapsed.avg.per_second',
        'smsp__average_warp_latency_issue_stalled_no_instruction', 'smsp__warps_eligible.avg.per_cycle_active',
        'smsp__issue_inst0.avg.pct_of_peak_sustained_active']
for k in keys:
    if k in hdr:
        i = hdr.index(k)
        print(f"{k} [{units[i]}]: {[r[i] for r in data]}")
stall = [(h, data[0][i]) for i, h in enumerate(hdr) if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
for h, v in sorted(stall, key=lambda x: -float(x[1] or 0))[:8]:
    print(h, v)
