#!/usr/bin/env python
"""profiles/ncu_current.json from an `ncu --set full` capture of the bench kernel: the figures bench.py quotes in
roofline.ncu / roofline.traffic (they are read from that file, never typed into bench.py).
usage: python tools/ncu_summary.py gpurun_out/<tag>_trace.ncu-rep <tag> [ray_steps_per_launch] > profiles/ncu_current.json"""
import csv, io, json, subprocess, sys
rep, tag = sys.argv[1], sys.argv[2]
steps = float(sys.argv[3]) if len(sys.argv) > 3 else 5.11e9
# a report, or the CSV of its raw page exported on the GPU box (gpurun brings back at most 64 MiB: reports stay there)
raw = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = [r for r in csv.reader(io.StringIO(raw)) if len(r) > 20]
hdr, units, vals = rows[0], rows[1], rows[2]
col = {h: i for i, h in enumerate(hdr)}


def get(name, scale=1.0):
    v = float(vals[col[name]])
    u = units[col[name]]
    mult = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(u, 1.0)
    return v * mult * scale


head = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
out = {
    "source": f"profiles/{tag}_trace_details.txt, profiles/{tag}_trace_raw_selected.txt (ncu --set full --clock-control none, one launch of `python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e`)",
    "commit": head, "kernel": vals[col["Kernel Name"]] if "Kernel Name" in col else None,
    "kernel_ms_under_ncu": get("gpu__time_duration.sum") * (1e-6 if units[col["gpu__time_duration.sum"]] == "ns" else 1e-3 if units[col["gpu__time_duration.sum"]] == "us" else 1.0),
    "dram_bytes_per_launch": int(get("dram__bytes_read.sum") + get("dram__bytes_write.sum")),
    "dram_bytes_read": int(get("dram__bytes_read.sum")), "dram_bytes_written": int(get("dram__bytes_write.sum")),
    "issue_slots_busy": get("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100.0,
    "fma_pipe_busy": get("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active") / 100.0,
    "l1tex_throughput": get("l1tex__throughput.avg.pct_of_peak_sustained_active") / 100.0,
    "l1_hit_rate": get("l1tex__t_sector_hit_rate.pct") / 100.0,
    "warp_instructions_per_ray_step": get("smsp__inst_executed.sum") / (steps / 32.0),
    "registers_per_thread": int(get("launch__registers_per_thread")),
    "local_load_instructions": int(get("sass__inst_executed_local_loads")), "local_store_instructions": int(get("sass__inst_executed_local_stores")),
    "active_threads_per_warp": get("smsp__thread_inst_executed_per_inst_executed.ratio"),
}
print(json.dumps(out, indent=1))
