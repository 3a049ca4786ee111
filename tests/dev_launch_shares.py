"""Developer tool: per-kernel share of a frame from an ncu launch list
(`ncu --metrics gpu__time_duration.sum --clock-control none --csv` of `python bench.py --steps 2
--warmup 3`):  python tests/dev_launch_shares.py gpurun_out/launches_bench.csv <frames> > profiles/rNN_launch_shares.csv"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
frames = float(sys.argv[2])
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
acc = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v = v / 1000 if r[ui] == "ns" else v * 1000 if r[ui] == "ms" else v
    acc.setdefault(r[ki], []).append(v)
ours = {k: v for k, v in acc.items()
        if "<unnamed>::" in k and not any(w in k for w in ("photo", "post_", "at::"))}
total = sum(sum(v) for v in ours.values()) / frames
print("# share of the frame per kernel, from the ncu launch list of `python bench.py --steps 2 --warmup 3 "
      "--no-cpu-baseline` (gpu__time_duration.sum, --clock-control none); cold-cache, serialised launches")
print("kernel,launches_per_frame,avg_us,us_per_frame,share")
for k, v in ours.items():
    per = sum(v) / frames
    print(f'"{k[:70]}",{len(v) / frames:.1f},{sum(v) / len(v):.1f},{per:.1f},{per / total:.3f}')
print(f"total,,,{total:.1f},1.000")
