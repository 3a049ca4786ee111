mkdir -p gpurun_out
SPLAT_B200_EXTRA_NVCC_FLAGS="-DSPLAT_UNIT_SUBS=1" python -c "
import sys; sys.path.insert(0,'splat-loam_b200'); import build; build.build(force=True)" > /dev/null 2>&1
python -m pytest tests/test_gpu_raster.py tests/test_gpu_raster_more.py tests/test_gpu_abi_more.py -x -q 2>&1 | tail -3
python bench.py --steps 30 --warmup 5 2>/dev/null | tail -1 > gpurun_out/quick.json
python - <<PY
import json
d=json.load(open("gpurun_out/quick.json"))
print("4x4 units:", round(d["value"],1), round(d["ms_per_step"],4), {k:round(v*1000,1) for k,v in d["stage_ms"].items()})
PY
