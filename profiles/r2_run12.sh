#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest12.txt 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r2_pytest12.txt
python bench.py > gpurun_out/r2_bench12_1gpu.json 2> gpurun_out/r2_bench12_1gpu.err; echo "bench1 rc=$?"; tail -3 gpurun_out/r2_bench12_1gpu.err
