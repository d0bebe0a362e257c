#!/bin/bash
# round-2 profiling run (1 GPU): launch list + ncu full captures of the final kernels
mkdir -p gpurun_out
A="--steps 1 --warmup 1 --no-cpu-baseline --nq-sweep= --no-microbench --no-extra --check 0"
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --nq-sweep= --no-microbench --no-extra > gpurun_out/plain_a.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2_launches_sparse.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --nq-sweep= --no-microbench --no-extra > gpurun_out/ncu_a.log 2>&1; echo "launch list rc=$?"
python bench.py $A > gpurun_out/plain_b.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:sparse_score_topk -s 1 -c 1 -f -o gpurun_out/r2_ncu_sparse python bench.py $A > gpurun_out/ncu_b.log 2>&1; echo "ncu sparse rc=$?"
python profiles/probe_small.py 4 100 3 > gpurun_out/plain_c.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'sparse_score_topk|topk_merge' -s 4 -c 2 -f -o gpurun_out/r2_ncu_sparse_nq4 python profiles/probe_small.py 4 100 3 > gpurun_out/ncu_c.log 2>&1; echo "ncu nq4 rc=$?"
python profiles/probe_dense_c5.py 1250000 1024 2 > gpurun_out/r2_dense_c5_probe.txt 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:dense_tc_kernel -s 1 -c 1 -f -o gpurun_out/r2_ncu_dense_c5 python profiles/probe_dense_c5.py 1250000 1024 1 > gpurun_out/ncu_d.log 2>&1; echo "ncu dense rc=$?"
cat gpurun_out/r2_dense_c5_probe.txt
