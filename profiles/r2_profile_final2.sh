#!/bin/bash
mkdir -p gpurun_out
A="--steps 1 --warmup 1 --no-cpu-baseline --nq-sweep= --no-microbench --no-extra --check 0"
python bench.py $A > gpurun_out/plain_b.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:sparse_score_topk -s 1 -c 1 -f -o gpurun_out/r2_ncu_sparse python bench.py $A > gpurun_out/ncu_b.log 2>&1; echo "ncu sparse rc=$?"
