#!/bin/bash
# 1 GPU: the no-alloc entry point's test alone in a fresh process, then the whole GPU suite (that test last), then config 5's
# per-GPU shard on one GPU (dense step: fused pass vs whole step, uncertified queries summed over the batches)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2_zz_alone.txt 2>&1; echo "zz alone rc=$?"; tail -3 gpurun_out/r2_zz_alone.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_1gpu.txt 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_pytest_gpu_1gpu.txt
timeout 900 python bench.py --gpus 1 --workload config5 --steps 5 --warmup 3 > gpurun_out/r2_c5_1gpu.json 2> gpurun_out/r2_c5_1gpu.err; echo "c5 rc=$?"
tail -2 gpurun_out/r2_c5_1gpu.err; cat gpurun_out/r2_c5_1gpu.json | head -c 3000
