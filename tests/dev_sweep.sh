mkdir -p gpurun_out
for cfg in "64 128" "32 128" "64 256" "96 256" "64 64"; do
set -- $cfg
SPLAT_B200_EXTRA_NVCC_FLAGS="-DCOMPOSITE_ENT_CAP=$1 -DCOMPOSITE_GATHER_PAIRS=$2" python -c "
import sys; sys.path.insert(0,'splat-loam_b200'); import build; build.build(force=True)" > /dev/null 2>&1
python bench.py --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/sweep.json
python - <<PY
import json
d=json.load(open("gpurun_out/sweep.json"))
print("ENT_CAP $1 PAIRS $2:", round(d["value"],1), {k:round(v*1000,1) for k,v in d["stage_ms"].items() if k.startswith("render")})
PY
done
