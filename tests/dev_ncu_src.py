"""Developer tool: summarise an ncu report's source page by code region (plateaus of the
per-warp execution count) with each region's share of instructions and stall samples.

    python tests/dev_ncu_src.py gpurun_out/x.ncu-rep [nwarps] [--full]
"""
import csv
import subprocess
import sys

rep = sys.argv[1]
nw = float(sys.argv[2]) if len(sys.argv) > 2 and not sys.argv[2].startswith("-") else 1.0
full = "--full" in sys.argv
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ci = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
for r in rows[2:]:
    try:
        data.append((int(r[ci["Instructions Executed"]]), r[ci["Source"]].strip(), int(r[ci["# Samples"]]),
                     float(r[ci["Avg. Threads Executed"]] or 0), {k: int(r[ci[k]] or 0) for k in stalls}))
    except Exception:
        pass
tot = sum(d[0] for d in data)
ts = sum(d[2] for d in data)
print("total inst", tot, "samples", ts)
if full:
    for i, (n, s, sm, thr, st) in enumerate(data):
        top = sorted(st.items(), key=lambda kv: -kv[1])[:2]
        print(f"{i:4d} {n / nw:9.1f} smp {100 * sm / ts:5.2f}% thr {thr:4.1f} {s[:64]:64s} {top}")
    sys.exit(0)
segs, cur = [], None
for i, (n, s, sm, thr, st) in enumerate(data):
    lvl = n / nw
    if cur is None or abs(lvl - cur["lvl"]) > 0.12 * max(cur["lvl"], 1):
        cur = {"lvl": lvl, "start": i, "n": 0, "inst": 0, "smp": 0, "first": s, "thr": 0, "st": {}}
        segs.append(cur)
    cur["n"] += 1; cur["inst"] += n; cur["smp"] += sm; cur["thr"] += thr * n
    for k, v in st.items():
        cur["st"][k] = cur["st"].get(k, 0) + v
for sg in segs:
    if sg["inst"] / tot > 0.003 or sg["smp"] / ts > 0.01:
        top = sorted(sg["st"].items(), key=lambda kv: -kv[1])[:3]
        print(f"idx {sg['start']:4d} len {sg['n']:3d} exec/warp {sg['lvl']:8.1f} inst {100 * sg['inst'] / tot:5.1f}% "
              f"samples {100 * sg['smp'] / ts:5.1f}% thr {sg['thr'] / max(sg['inst'], 1):4.1f} {top}")
