mkdir -p gpurun_out
python -m pytest tests/test_gpu_raster.py tests/test_gpu_raster_more.py tests/test_gpu_knn.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -4
python bench.py --steps 30 --warmup 5 2>/dev/null | tail -1 > gpurun_out/ab_tile.json
SPLAT_B200_SORT=global python bench.py --steps 30 --warmup 5 2>/dev/null | tail -1 > gpurun_out/ab_global.json
python - <<PY
import json
for n in ("tile","global"):
    d=json.load(open(f"gpurun_out/ab_{n}.json"))
    print(n, round(d["value"],1), round(d["ms_per_step"],4), {k:round(v*1000,1) for k,v in d["stage_ms"].items()})
PY
