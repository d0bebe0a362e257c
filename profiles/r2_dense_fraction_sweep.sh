#!/bin/bash
# config 4 on one GPU: the share of docs above which a term is stored as an impact row (MSR_SPARSE_DENSE_FRACTION)
mkdir -p gpurun_out
B="--steps 10 --warmup 3 --no-cpu-baseline --nq-sweep= --no-microbench --no-extra"
for f in 0.15 0.06 0.08 0.11 0.20 0.15; do
  MSR_SPARSE_DENSE_FRACTION=$f timeout 300 python bench.py $B > gpurun_out/r2_df_$f.json 2> gpurun_out/r2_df_$f.err
  python - "$f" <<'PY'
import json, sys
f = sys.argv[1]
d = json.loads(open(f"gpurun_out/r2_df_{f}.json").read().strip().splitlines()[-1])
print(f"dense fraction {f}: {d['value']:.0f} q/s, {d['ms_per_step']:.3f} ms/step, scorer {d['roofline']['kernel_ms_per_launch']:.3f} ms")
PY
done | tee gpurun_out/r2_dense_fraction_sweep.txt
