#!/bin/bash
mkdir -p gpurun_out
python profiles/probe_quantise.py 3 > gpurun_out/plain_q.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:topk_rows_kernel -s 2 -c 1 -f -o gpurun_out/r2_ncu_rows python profiles/probe_quantise.py 3 > gpurun_out/ncu_q.log 2>&1; echo "ncu rows rc=$?"
