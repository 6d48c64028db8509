#!/bin/bash
# usage: retry.sh <timeout-seconds> <job script> [gpus]  -- retries gpurun while the pod answers busy (rc 3)
T=$1; J=$2; G=${3:-1}
for i in $(seq 1 40); do
  if [ "$G" -gt 1 ]; then /usr/local/graft/bin/gpurun --gpus $G --timeout $T -- "bash $J" > gpurun_out/.retry_$(basename $J).log 2>&1; rc=$?
  else /usr/local/graft/bin/gpurun --timeout $T -- "bash $J" > gpurun_out/.retry_$(basename $J).log 2>&1; rc=$?; fi
  if grep -q "status=transient" gpurun_out/.retry_$(basename $J).log; then sleep 60; continue; fi
  break
done
tail -40 gpurun_out/.retry_$(basename $J).log
