#!/bin/bash
# Dev helper (this container): run scripts/_tmp_run.sh on a B200 through gpurun, retrying while
# the pod answers "busy" (exit 3).  usage: scripts/gpu_call.sh TAG [TIMEOUT_S] [GPUS]
TAG=${1:-x}; TMO=${2:-1500}; G=${3:-1}
for try in $(seq 1 30); do
  if [ "$G" = "1" ]; then
    /usr/local/graft/bin/gpurun --timeout $TMO -- 'bash scripts/_tmp_run.sh' > gpurun_out/call_$TAG.log 2>&1
  else
    /usr/local/graft/bin/gpurun --gpus $G --timeout $TMO -- 'bash scripts/_tmp_run.sh' > gpurun_out/call_$TAG.log 2>&1
  fi
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (try $try)"; exit $rc; fi
  sleep 90
done
echo "gave up"; exit 3
