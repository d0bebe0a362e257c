#!/bin/bash
# final build at N = 4 and N = 2 (20 steps), launched as the driver launches it
mkdir -p gpurun_out
for n in 4 2; do
  MSR_PX_PROFILE=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2957$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_bench21_${n}gpu.json 2> gpurun_out/r2_bench21_${n}gpu.err; echo "bench$n rc=$?"
done
python - <<'PY'
import json
for n in (4, 2):
    d = json.loads(open(f"gpurun_out/r2_bench21_{n}gpu.json").read().strip().splitlines()[-1])
    print(n, d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("parity_check"), json.dumps(d.get("exchange_phases_ms"))[:200])
PY
