#!/bin/bash
# usage: scratch/run_variants_n.sh <n_docs> "<lib or ->:<tile_docs>" ...
n=$1; shift
for spec in "$@"; do
  lib=${spec%%:*}; td=${spec##*:}
  if [ "$lib" != "-" ]; then export MSR_LIB_PATH=$PWD/scratch/libmsr_$lib.so; else unset MSR_LIB_PATH; fi
  python bench.py --n-docs $n --tile-docs $td --no-cpu-baseline --steps 5 --warmup 3 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$n $spec', round(j['value']), round(j['roofline']['kernel_ms_per_launch'],3), round(j['roofline']['merge_ms_per_step'],3), j['index']['n_tiles'], j['parity_check'])"
done
