#!/bin/bash
# final build on one GPU: the whole GPU suite, then the driver's default command
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_1gpu_final.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu_1gpu_final.txt
timeout 400 python bench.py > gpurun_out/r2_bench22_1gpu.json 2> gpurun_out/r2_bench22_1gpu.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2_bench22_1gpu.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["roofline"]["frac"], d["roofline"].get("traffic"))
print(json.dumps(d.get("query_construction"))[:200])
print(d["extra"]["dense"]["value"], d["extra"]["hybrid"]["value"], d["cpu_baseline"]["value"], d["clocks"])
PY
