#!/bin/bash
mkdir -p gpurun_out
for N in 8 4; do
MSR_PX_PROFILE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2956$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench14_${N}gpu.json 2> gpurun_out/r2_bench14_${N}gpu.err; echo "bench$N rc=$?"
done
