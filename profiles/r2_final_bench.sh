#!/bin/bash
# final build on one GPU: the whole GPU suite, then the driver's two commands
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_1gpu_final.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu_1gpu_final.txt
timeout 900 python bench.py > gpurun_out/r2_bench20_1gpu.json 2> gpurun_out/r2_bench20_1gpu.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/r2_bench20_reference.json 2> gpurun_out/r2_bench20_reference.err; echo "reference rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2_bench20_1gpu.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["roofline"]["frac"], d["roofline"].get("traffic"), json.dumps(d["roofline"].get("binding"))[:600])
print(json.dumps(d.get("query_construction"))[:400])
print(json.dumps(d["extra"]["dense"])[:300])
r = json.loads(open("gpurun_out/r2_bench20_reference.json").read().strip().splitlines()[-1])
print("reference", r["value"], r["cpu_baseline"])
PY
