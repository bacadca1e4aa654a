#!/bin/bash
# tools/gpu_variants.sh NAME... : time the fixed profiling workload with each lib/variants/libmgb_NAME.so (A/B builds)
for v in "$@"; do
  MGB_LIB=$PWD/magritte_b200/lib/variants/libmgb_$v.so python tools/prof_solve.py --npoints 30000 --nside 2 --reps 2 > gpurun_out/pv_$v.log 2>&1
  echo -n "$v: "; tail -1 gpurun_out/pv_$v.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_solve', d['ms_solve'], 'J-less check entries', d['n_entries'])" || tail -3 gpurun_out/pv_$v.log
done
