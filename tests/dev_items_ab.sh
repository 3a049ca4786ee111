mkdir -p gpurun_out
for it in 8 12 16; do
SPLAT_B200_SORT_ITEMS=$it python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/it.json
python - <<PY
import json
d=json.load(open("gpurun_out/it.json"))
print("items $it:", round(d["value"],1), {k:round(v*1000,1) for k,v in d["stage_ms"].items() if k in ("emit","sort")})
PY
done
