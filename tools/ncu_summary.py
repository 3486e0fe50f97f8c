"""Summarise an ncu --set full report (read here, no GPU needed) into a small JSON for profiles/.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep "command line that was profiled" > profiles/rNN_ncu_full_summary.json"""
import csv
import json
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.avg"]

rep, cmd = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
head, units = rows[0], rows[1]
out = {"command": cmd, "kernels": []}
for r in rows[2:]:
    k = {"Kernel Name": r[head.index("Kernel Name")], "Grid Size": r[head.index("Grid Size")],
         "Block Size": r[head.index("Block Size")]}
    for name in KEYS:
        if name in head:
            i = head.index(name)
            k[name] = f"{r[i]} {units[i]}".strip()
    out["kernels"].append(k)
print(json.dumps(out, indent=1))
