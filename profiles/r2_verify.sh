#!/bin/bash
# verification run (2 GPUs): whole GPU suite, default 1-GPU bench line, 2-GPU bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_verify.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_pytest_verify.txt
python bench.py > gpurun_out/r2_bench_verify_1gpu.json 2> gpurun_out/r2_bench_verify_1gpu.err; echo "bench1 rc=$?"; tail -3 gpurun_out/r2_bench_verify_1gpu.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_bench_verify_2gpu.json 2> gpurun_out/r2_bench_verify_2gpu.err; echo "bench2 rc=$?"
python __graft_entry__.py smoke 2>&1 | tail -2
