#!/bin/bash
# reduced scaling run on an 8-GPU box: N = 8 (with the BASELINE windows), 4, 2 (no windows), 1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 100 > gpurun_out/r2_scale8.json 2> gpurun_out/r2_scale8.err
for n in 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 100 --no-window > gpurun_out/r2_scale$n.json 2> gpurun_out/r2_scale$n.err
done
python bench.py --steps 100 --no-cpu-baseline > gpurun_out/r2_scale1.json 2>/dev/null
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --impl reference --gpus 8 --steps 30 --no-cpu-baseline > gpurun_out/r2_scale8_ref.json 2>/dev/null
echo done
