#!/bin/bash
mkdir -p gpurun_out
MSR_PX_PROFILE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29548 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench7_8gpu.json 2> gpurun_out/r2_bench7_8gpu.err; echo "bench8 rc=$?"
timeout 300 python -m pytest tests/test_peer_exchange_gpu.py -m gpu -x -q > gpurun_out/r2_pytest7.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest7.txt
