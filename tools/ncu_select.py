#!/usr/bin/env python
"""The metrics the profiles/ summaries quote, from `ncu -i rep --page raw --csv` output (one kernel).
usage: python tools/ncu_select.py gpurun_out/<tag>_raw.csv "<what was captured>" > profiles/<tag>_raw_selected.txt"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
rows = [r for r in rows if len(r) > 20]
hdr, units, vals = rows[0], rows[1], rows[2]
col = {h: i for i, h in enumerate(hdr)}
print("# ncu --set full --clock-control none: %s" % (sys.argv[2] if len(sys.argv) > 2 else ""))
if "Kernel Name" in col: print("# kernel: %s  grid %s block %s" % (vals[col["Kernel Name"]], vals[col.get("Grid Size", 0)], vals[col.get("Block Size", 0)]))
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__shared_mem_per_block_static", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_global_ld.sum",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores", "smsp__thread_inst_executed_per_inst_executed.ratio"]
for k in want:
    if k in col: print("%s [%s] = %s" % (k, units[col[k]], vals[col[k]]))
for h in hdr:
    if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
        try:
            if float(vals[col[h]]) >= 0.2: print("%s = %s" % (h, vals[col[h]]))
        except ValueError: pass
