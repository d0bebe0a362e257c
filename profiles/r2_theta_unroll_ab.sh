#!/bin/bash
# 1 GPU: A/B of MSR_THETA_UNROLL (1 = default build, 2, 4) on config 4, two interleaved rounds; then the driver's own two
# commands (default bench line, reference arm)
mkdir -p gpurun_out
B="--steps 20 --warmup 5 --no-cpu-baseline --nq-sweep= --no-microbench --no-extra"
for round in a b; do
  for v in base tu2 tu4; do
    if [ $v = base ]; then unset MSR_LIB_PATH; else export MSR_LIB_PATH=$PWD/mllm_sparse_retrieval_b200/libmsr_$v.so; fi
    timeout 300 python bench.py $B > gpurun_out/r2_tu_${v}_${round}.json 2> gpurun_out/r2_tu_${v}_${round}.err; echo "$v $round rc=$?"
  done
done
unset MSR_LIB_PATH
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2_tu_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["value"]), d["ms_per_step"], d["roofline"]["kernel_ms_per_launch"])
    except Exception as e:
        print(f, "unreadable", e)
PY
timeout 900 python bench.py > gpurun_out/r2_bench18_1gpu.json 2> gpurun_out/r2_bench18_1gpu.err; echo "default bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/r2_bench18_reference.json 2> gpurun_out/r2_bench18_reference.err; echo "reference arm rc=$?"
head -c 600 gpurun_out/r2_bench18_1gpu.json; echo; head -c 600 gpurun_out/r2_bench18_reference.json
