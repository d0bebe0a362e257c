#!/bin/bash
# round-2 GPU run 3 (2 GPUs): all GPU tests incl. the 2-GPU exchange tests, then 2-GPU bench with the fused exchange
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest3.txt 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r2_pytest3.txt
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench3_1gpu.json 2> gpurun_out/r2_bench3_1gpu.err; echo "bench1 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_bench3_2gpu.json 2> gpurun_out/r2_bench3_2gpu.err; echo "bench2 rc=$?"
tail -3 gpurun_out/r2_bench3_2gpu.err
