#!/bin/bash
# tools/gpu_check.sh TAG : on the GPU box -- parity tests, then the fixed profiling workload (tools/prof_solve.py); logs in gpurun_out/
TAG=${1:-x}
python -m pytest tests -m gpu -x -q > gpurun_out/t_$TAG.log 2>&1; tail -3 gpurun_out/t_$TAG.log
python tools/prof_solve.py --npoints 30000 --nside 2 --reps 2 > gpurun_out/p_$TAG.log 2>&1; tail -1 gpurun_out/p_$TAG.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_solve', d['ms_solve'])"
