#!/bin/bash
# usage: scratch/run_variants.sh "<lib or ->:<tile_docs>" ...
for spec in "$@"; do
  lib=${spec%%:*}; td=${spec##*:}
  if [ "$lib" != "-" ]; then export MSR_LIB_PATH=$PWD/scratch/libmsr_$lib.so; else unset MSR_LIB_PATH; fi
  python bench.py --tile-docs $td --no-cpu-baseline --steps 5 --warmup 3 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('$spec', round(j['value']), round(j['roofline']['kernel_ms_per_launch'],3), j['roofline']['merge_ms_per_step'], j['index']['n_tiles'], j['parity_check'])"
done
