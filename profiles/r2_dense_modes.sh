#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_topk_fuse_dense_gpu.py tests/test_zz_no_alloc.py tests/test_host_layer_gpu.py -m gpu -x -q 2>&1 | tail -3
for a in "1250000 4096" "80000 4096" "25000 4096" "1000000 768"; do timeout 300 python profiles/probe_dense_modes.py $a 2>&1 | tail -4; done | tee gpurun_out/r2_dense_modes_after.txt
