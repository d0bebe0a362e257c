#!/bin/bash
mkdir -p gpurun_out
python profiles/probe_dense_c5.py 1250000 1024 2 > gpurun_out/plain_d.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_dense_c5.csv python profiles/probe_dense_c5.py 1250000 1024 1 > gpurun_out/ncu_d.log 2>&1; echo "rc=$?"
cat gpurun_out/plain_d.log | tail -2
python - <<'PY'
import csv
rows = [r for r in csv.reader(open("gpurun_out/r2_launches_dense_c5.csv")) if len(r) > 5]
hdr = rows[0]; ik = hdr.index("Kernel Name"); iv = hdr.index("Metric Value")
# the last search = the launches after the last to_f16_kernel
names = [(r[ik][:60], float(r[iv].replace(",", ""))) for r in rows[1:]]
last = max(i for i, (n, _) in enumerate(names) if "to_f16" in n)
for n, v in names[last:]:
    print(f"{v/1e3:10.1f} us  {n}")
PY
