#!/bin/bash
# round-2 GPU run 4 (8 GPUs): scaling bench 8 / 4 with the shared threshold + fused barrier
mkdir -p gpurun_out
for N in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench4_${N}gpu.json 2> gpurun_out/r2_bench4_${N}gpu.err; echo "bench$N rc=$?"
done
MSR_SPARSE_NO_PEER_THETA=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench4_8gpu_notheta.json 2> gpurun_out/r2_bench4_8gpu_notheta.err; echo "bench8 notheta rc=$?"
