#!/bin/bash
mkdir -p gpurun_out
MSR_PX_PROFILE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench5_8gpu_phases.json 2> gpurun_out/r2_bench5_8gpu.err; echo "rc=$?"
