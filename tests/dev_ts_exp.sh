for e in 0 1 2; do
SPLAT_B200_EXTRA_NVCC_FLAGS="-DSORT_TRACE -DTS_EXP=$e" python -c "
import sys; sys.path.insert(0,'splat-loam_b200'); import build; build.build(force=True)" > /dev/null 2>&1
echo "== TS_EXP=$e"; python tests/dev_tile_sort_trace.py | tail -2
done
