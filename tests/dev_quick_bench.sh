mkdir -p gpurun_out
python -m pytest tests/test_gpu_raster.py tests/test_gpu_raster_more.py tests/test_gpu_abi_more.py -x -q 2>&1 | tail -3
python bench.py --steps 30 --warmup 5 2>/dev/null | tail -1 > gpurun_out/quick.json
python - <<PY
import json
d=json.load(open("gpurun_out/quick.json"))
print(round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), {k:round(v*1000,1) for k,v in d["stage_ms"].items()})
PY
