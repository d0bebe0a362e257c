#!/bin/bash
# final-source recapture (1 GPU): ncu --set full of one scorer launch of the default bench + launch list
mkdir -p gpurun_out
A="--steps 1 --warmup 1 --no-cpu-baseline --nq-sweep= --no-microbench --no-extra --check 0"
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --nq-sweep= --no-microbench --no-extra > gpurun_out/plain_a.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2_launches_sparse.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --nq-sweep= --no-microbench --no-extra > gpurun_out/ncu_a.log 2>&1; echo "launch list rc=$?"
