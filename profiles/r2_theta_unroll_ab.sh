#!/bin/bash
# A/B on one box: independent accumulator loads per step of the threshold pass (MSR_THETA_UNROLL = 1 / 2 / 4), config 4
mkdir -p gpurun_out
A="--steps 10 --warmup 3 --no-cpu-baseline --nq-sweep= --no-microbench --no-extra"
for L in libmsr libmsr_tu2 libmsr_tu4 libmsr; do
MSR_LIB_PATH=$PWD/mllm_sparse_retrieval_b200/$L.so python bench.py $A > gpurun_out/r2_ab_$L.json 2> gpurun_out/r2_ab_$L.err; echo "$L rc=$?"
python - <<PY
import json
r=json.load(open('gpurun_out/r2_ab_$L.json')); print('$L', round(r['value']), r['roofline']['kernel_ms_per_launch'], r['parity_check'][:30])
PY
done
for L in libmsr libmsr_tu2 libmsr_tu4; do
MSR_LIB_PATH=$PWD/mllm_sparse_retrieval_b200/$L.so python bench.py --n-docs 125000 $A > gpurun_out/r2_ab125_$L.json 2>> gpurun_out/r2_ab_$L.err
python - <<PY
import json
r=json.load(open('gpurun_out/r2_ab125_$L.json')); print('125k $L', round(r['value']), r['roofline']['kernel_ms_per_launch'], r['parity_check'][:30])
PY
done
