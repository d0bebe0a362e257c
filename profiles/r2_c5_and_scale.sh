#!/bin/bash
# 8 GPUs: default bench at N=8 (20 steps) + config 5 (10 M docs sparse + 4096-d dense, weak scaling, 10 steps)
mkdir -p gpurun_out
MSR_PX_PROFILE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29578 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench17_8gpu.json 2> gpurun_out/r2_bench17_8gpu.err; echo "bench8 rc=$?"
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29579 bench.py --gpus 8 --workload config5 --steps 10 --warmup 3 > gpurun_out/r2_c5_8gpu.json 2> gpurun_out/r2_c5_8gpu.err; echo "c5 rc=$?"
tail -3 gpurun_out/r2_c5_8gpu.err
