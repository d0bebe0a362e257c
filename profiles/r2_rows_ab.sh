#!/bin/bash
mkdir -p gpurun_out
for v in base rf4 base rf4; do
  if [ $v = base ]; then unset MSR_LIB_PATH; else export MSR_LIB_PATH=$PWD/mllm_sparse_retrieval_b200/libmsr_$v.so; fi
  echo "== $v"; timeout 300 python profiles/probe_rows.py 2>&1 | tail -4
done | tee gpurun_out/r2_rows_ab.txt
unset MSR_LIB_PATH
timeout 600 python -m pytest tests/test_topk_fuse_dense_gpu.py -m gpu -x -q 2>&1 | tail -3
