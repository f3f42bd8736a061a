for z in 0 1; do
  if [ $z = 1 ]; then export UMRS_ZERO_COPY=1; fi
  for rep in 1 2; do
  timeout 300 python bench.py --no-cpu-baseline --steps 300 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('zero_copy=$z bench: %.4e rays/s  %.4f ms/step  e2e %.4e (%.4f ms)' % (d['value'],d['ms_per_step'],d['e2e']['value'],d['e2e']['ms_per_step']))"
  done
done > gpurun_out/zc.log 2>&1
