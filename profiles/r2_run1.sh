#!/bin/bash
# round-2 GPU run 1: sanity tests, default bench with the new lines, shard-shaped bench, ncu full captures (1 GPU)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest1.txt 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r2_pytest1.txt
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err; echo "bench rc=$?"
python bench.py --n-docs 125000 --steps 10 --warmup 3 --no-cpu-baseline --nq-sweep "" > gpurun_out/r2_bench_125k.json 2> gpurun_out/r2_bench_125k.err; echo "bench125k rc=$?"
A="--steps 1 --warmup 1 --no-cpu-baseline --nq-sweep= --no-microbench --check 0"
python bench.py $A > gpurun_out/plain1.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:sparse_score_topk -s 1 -c 1 -f -o gpurun_out/r2_ncu_sparse_base python bench.py $A > gpurun_out/ncu1.log 2>&1
echo "ncu1 rc=$?"
python bench.py --n-docs 125000 $A > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:sparse_score_topk -s 1 -c 1 -f -o gpurun_out/r2_ncu_sparse_base_125k python bench.py --n-docs 125000 $A > gpurun_out/ncu2.log 2>&1
echo "ncu2 rc=$?"
ls -la gpurun_out
