#!/bin/bash
# scaling runs on an 8-GPU box: N = 8, 4, 2, 1 of our arm, the reference arm and the NCCL variant at 8
mkdir -p gpurun_out
python -m pytest tests/test_gpu_exchange_mp.py tests/test_gpu_abi_more.py -m gpu -q 2>&1 | tail -2
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 100 > gpurun_out/r2_scale$n.json 2> gpurun_out/r2_scale$n.err
  grep -c "^{" gpurun_out/r2_scale$n.json
done
python bench.py --steps 100 --no-cpu-baseline > gpurun_out/r2_scale1.json 2>/dev/null
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --impl reference --gpus 8 --steps 50 --no-cpu-baseline > gpurun_out/r2_scale8_ref.json 2>/dev/null
SPLAT_B200_XCHG=nccl python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 100 --no-window > gpurun_out/r2_scale8_nccl.json 2>/dev/null
echo done
