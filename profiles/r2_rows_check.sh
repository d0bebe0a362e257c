#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_topk_fuse_dense_gpu.py tests/test_cli_gpu.py tests/test_zz_no_alloc.py tests/test_host_layer_gpu.py -m gpu -x -q > gpurun_out/r2_pytest_rows.txt 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2_pytest_rows.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --nq-sweep= --no-microbench > gpurun_out/r2_bench19_rows.json 2> gpurun_out/r2_bench19_rows.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2_bench19_rows.json").read().strip().splitlines()[-1])
print(d["value"], json.dumps(d.get("query_construction"))[:700])
print(json.dumps(d.get("extra"))[:1500])
PY
