#!/bin/bash
# round-2 GPU run 2: microbenchmarks (incl. legacy IMMA) + small-batch launch lists
mkdir -p gpurun_out
python - > gpurun_out/r2_microbench.txt 2>&1 <<'PY'
import sys; sys.path.insert(0,'.')
from mllm_sparse_retrieval_b200 import ops
names=["L2_READ","HBM_READ","RED_SHARED","ISSUE_DP4A","ALU_PRMT","LDS_WAVEFRONTS","IMMA","IMMA_PRMT8"]
for i,n in enumerate(names):
    b = (48<<20) if i==0 else (4<<30) if i==1 else 0
    v = ops.microbench(i, b, 5, 0)
    print(f"{n:16s} {v:.4e} per s   per SM-clock(1.965GHz,148): {v/148/1.965e9:.3f}")
PY
cat gpurun_out/r2_microbench.txt
for K in 100 1000; do
python profiles/probe_small.py 4 $K 3 > gpurun_out/probe_$K.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'sparse_score|topk_merge|query_prep' --csv --log-file gpurun_out/r2_launches_small_nq4_k$K.csv python profiles/probe_small.py 4 $K 3 > gpurun_out/ncu_small_$K.log 2>&1
echo "k=$K rc=$?"
done
