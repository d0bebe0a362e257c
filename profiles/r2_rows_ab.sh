#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_topk_fuse_dense_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
for v in base mb4 mb2 base; do
  if [ $v = base ]; then unset MSR_LIB_PATH; else export MSR_LIB_PATH=$PWD/mllm_sparse_retrieval_b200/libmsr_$v.so; fi
  echo "== $v"; timeout 300 python profiles/probe_rows.py 2>&1 | tail -4
done | tee gpurun_out/r2_rows_ab2.txt
