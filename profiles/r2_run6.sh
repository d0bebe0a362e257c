#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_peer_exchange_gpu.py tests/test_sparse_gpu.py -m gpu -x -q > gpurun_out/r2_pytest6.txt 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_pytest6.txt
for N in 8 4 2; do
MSR_PX_PROFILE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench6_${N}gpu.json 2> gpurun_out/r2_bench6_${N}gpu.err; echo "bench$N rc=$?"
done
