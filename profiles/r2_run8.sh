#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_peer_exchange_gpu.py tests/test_sparse_gpu.py -m gpu -x -q > gpurun_out/r2_pytest8.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest8.txt
MSR_PX_PROFILE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29558 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench8_2gpu.json 2> gpurun_out/r2_bench8_2gpu.err; echo "bench2 rc=$?"
