"""Developer tool: one block of headline metrics per kernel of an ncu report
(`ncu --set full`), the last launch of each kernel name:

    python tests/dev_ncu_summary.py gpurun_out/x.ncu-rep > profiles/rNN_kernels_ncu.txt
"""
import csv
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_issued.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "nvltx__bytes.sum", "nvlrx__bytes.sum",
]

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True,
                     text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
last = {}
for r in rows[2:]:
    last[r[hdr.index("Kernel Name")]] = r
print(f"# {sys.argv[1]}: ncu --set full --clock-control none, last launch of each kernel")
for name, r in last.items():
    print(f"\n{name[:110]}")
    for m in METRICS:
        if m in hdr:
            i = hdr.index(m)
            print(f"    {m:62s} {r[i]:>16s} {units[i]}")
