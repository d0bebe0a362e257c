#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_topk_fuse_dense_gpu.py tests/test_zz_no_alloc.py tests/test_cli_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python profiles/probe_rows.py 2>&1 | tail -4 | tee gpurun_out/r2_rows_ab3.txt
