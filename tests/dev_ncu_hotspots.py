"""Developer tool: source-level hot-spot table of a kernel from an ncu report captured with
`--set full --import-source on` (the library is built with -lineinfo): the top source lines by
executed warp instructions, with their share of the stall samples and the two main stall reasons.

    python tests/dev_ncu_hotspots.py gpurun_out/x.ncu-rep render_forward [top]
"""
import csv
import subprocess
import sys

rep, which = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 20
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
secs, cur = [], None
for r in csv.reader(out.splitlines()):
    if r and r[0] == "File Path":
        cur = {"file": r[1], "rows": [], "hdr": None, "fn": ""}
        secs.append(cur)
    elif r and r[0] == "Function Name" and cur is not None:
        cur["fn"] = r[1]
    elif r and r[0] == "Line No" and cur is not None:
        cur["hdr"] = r
    elif cur and cur["hdr"] and r and r[0].isdigit():
        cur["rows"].append(r)


def num(x):
    try:
        return int(x)
    except ValueError:
        return 0


lines, stall_tot = [], {}
for s in secs:
    if which not in s["fn"]:
        continue
    ci = {}
    for i, k in enumerate(s["hdr"]):
        ci.setdefault(k, i)
    stalls = [k for k in s["hdr"] if k.startswith("stall_") and "Not Issued" not in k]
    for r in s["rows"]:
        st = {k[6:]: num(r[ci[k]]) for k in stalls}
        for k, v in st.items():
            stall_tot[k] = stall_tot.get(k, 0) + v
        lines.append((num(r[ci["Instructions Executed"]]), num(r[ci["# Samples"]]),
                      s["file"].split("/")[-1], r[0], r[1].strip(), st))
merged = {}          # several launches of the kernel in one report: one row per source line
for inst, smp, f, ln, src, st in lines:
    m = merged.setdefault((f, ln), [0, 0, f, ln, src, {}])
    m[0] += inst
    m[1] += smp
    for k, v in st.items():
        m[5][k] = m[5].get(k, 0) + v
lines = [tuple(m) for m in merged.values()]
ti, ts = sum(x[0] for x in lines), sum(x[1] for x in lines)
print(f"# {rep}: {which}: {ti} warp instructions, {ts} stall samples")
print("# stall reasons (share of samples): " +
      ", ".join(f"{k} {100 * v / max(ts, 1):.1f}%" for k, v in sorted(stall_tot.items(), key=lambda kv: -kv[1])[:7]))
print(f"# {'file:line':28s} {'inst %':>7s} {'smp %':>7s}  source  [main stalls]")
for inst, smp, f, ln, src, st in sorted(lines, key=lambda x: -x[0])[:top]:
    main = sorted(st.items(), key=lambda kv: -kv[1])[:2]
    print(f"{f + ':' + ln:30s} {100 * inst / ti:6.1f}% {100 * smp / max(ts, 1):6.1f}%  {src[:78]:78s} "
          f"{[(k, v) for k, v in main if v]}")
