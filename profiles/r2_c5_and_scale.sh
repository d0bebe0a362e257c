#!/bin/bash
# 8 GPUs: whole GPU suite (multi-GPU tests included), default bench at N=8 (20 steps), config 5 (10 M docs sparse + 4096-d
# dense, weak scaling, 10 steps)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_8gpu_box.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu_8gpu_box.txt
MSR_PX_PROFILE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29578 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench17_8gpu.json 2> gpurun_out/r2_bench17_8gpu.err; echo "bench8 rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29579 bench.py --gpus 8 --workload config5 --steps 10 --warmup 3 > gpurun_out/r2_c5_8gpu.json 2> gpurun_out/r2_c5_8gpu.err; echo "c5 rc=$?"
tail -3 gpurun_out/r2_c5_8gpu.err
python - <<'PY'
import json
for f in ("gpurun_out/r2_bench17_8gpu.json", "gpurun_out/r2_c5_8gpu.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"), json.dumps(d.get("dense")), json.dumps(d.get("exchange_phases_ms"))[:300])
    except Exception as e:
        print(f, "unreadable", e)
PY
