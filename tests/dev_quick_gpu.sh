#!/bin/bash
# quick GPU check: parity tests of the rasterizer + bench stage times
python -m pytest tests/test_gpu_raster.py tests/test_gpu_raster_more.py tests/test_gpu_exchange.py -m gpu -q 2>&1 | tail -3
python bench.py --steps 100 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith(chr(123))][-1]); print(round(d['value'],1), round(d['e2e']['value'],1), {k:round(v*1000,1) for k,v in d['stage_ms'].items()})
"
