mkdir -p gpurun_out
echo skip-tests
for pdl in 1 0; do
SPLAT_B200_PDL=$pdl timeout 120 python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/pdl.json
python - <<PY
import json
d=json.load(open("gpurun_out/pdl.json"))
print("PDL=$pdl:", round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), {k:round(v*1000,1) for k,v in d["stage_ms"].items()})
PY
done
