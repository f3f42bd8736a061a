#!/bin/bash
# Dev helper: GPU parity tests + same-box A/B timings of kernel variants.
# usage: scripts/gpu_ab.sh "VAR=val ..." "VAR=val ..."   (one quoted env set per variant: UMRS_LIB=build/x.so,
# UMRS_TUNE_CTA_SHAPE=3, UMRS_TUNE_CAP=..., see ut_multirobot_sim_b200/_lib.py)
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
i=0
for V in "$@"; do
  for rep in 1 2; do
  (
    export $V
    echo "== variant $i: $V (rep $rep)"
    timeout 300 python scripts/quick_gpu_check.py 4096 2>&1 | grep -v "^M=64"
    timeout 300 python bench.py --no-cpu-baseline --steps 300 2> /dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('bench: %.4e rays/s  %.4f ms/step  e2e %.4e  tests/ray %.2f' % (d['value'],d['ms_per_step'],d['e2e']['value'],d['roofline']['tests_per_ray']))"
    timeout 300 python scripts/bench_configs.py 2 4 2>/dev/null | grep '^{' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('config', d['config'], d.get('worlds',''), '%.4f ms  %.4e rays/s' % (d['ms_per_step'], d['rays_per_s']))"
  ) >> gpurun_out/ab.log 2>&1
  done
  i=$((i+1))
done
