mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/sort_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_sort.log 2>&1
python - <<PY
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/sort_launches.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); ui=hdr.index("Metric Unit")
acc=collections.defaultdict(list)
for r in rows[1:]:
    v=float(r[vi].replace(",","")); u=r[ui]
    if u=="ns": v/=1000
    elif u=="ms": v*=1000
    acc[r[ki][:60]].append(v)
for k,v in acc.items(): print("%-62s n=%3d avg %.1f us"%(k,len(v),sum(v)/len(v)))
PY
