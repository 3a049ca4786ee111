mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --steps 30 --warmup 5 2>/dev/null | tail -1 > gpurun_out/ab_tile.json
SPLAT_B200_SORT=global python bench.py --steps 30 --warmup 5 2>/dev/null | tail -1 > gpurun_out/ab_global.json
python - <<PY
import json
for n in ("tile","global"):
    d=json.load(open(f"gpurun_out/ab_{n}.json"))
    print(n, round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), {k:round(v*1000,1) for k,v in d["stage_ms"].items()})
PY
