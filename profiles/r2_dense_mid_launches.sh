#!/bin/bash
mkdir -p gpurun_out
python profiles/probe_dense_mid.py 64 > gpurun_out/plain_m.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_dense_mid.csv python profiles/probe_dense_mid.py 64 > gpurun_out/ncu_m.log 2>&1; echo "rc=$?"
python - <<'PY'
import csv
rows = [r for r in csv.reader(open("gpurun_out/r2_launches_dense_mid.csv")) if len(r) > 5]
hdr = rows[0]; ik = hdr.index("Kernel Name"); iv = hdr.index("Metric Value")
names = [(r[ik][:70], float(r[iv].replace(",", ""))) for r in rows[1:]]
last = max(i for i, (n, _) in enumerate(names) if "to_f16" in n)
tot = 0
for n, v in names[last:]:
    print(f"{v/1e3:10.1f} us  {n}"); tot += v
print(f"{tot/1e3:10.1f} us  total of kernels")
PY
