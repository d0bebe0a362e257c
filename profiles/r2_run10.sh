#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest10.txt 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2_pytest10.txt
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench10_1gpu.json 2> gpurun_out/r2_bench10_1gpu.err; echo "bench1 rc=$?"
