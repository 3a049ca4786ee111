#!/bin/bash
# the round's evidence in one GPU call: both bench arms, the ncu launch list of the bench
# command, one ncu --set full capture of every kernel of a frame
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err
python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python tests/dev_profile.py C2 3 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on --launch-skip 20 -o gpurun_out/kernels_full -f \
    python tests/dev_profile.py C2 3 > gpurun_out/prof_ncu.log 2>&1
tail -1 gpurun_out/bench_ours.json | cut -c1-600
tail -1 gpurun_out/bench_ref.json | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
