#!/bin/bash
# full GPU suite (multi-GPU tests need >= 2 GPUs) + smoke(), as the driver runs them at round end
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_final.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_pytest_gpu_final.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
