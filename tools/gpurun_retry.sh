#!/bin/bash
# tools/gpurun_retry.sh <timeout_s> <logfile> <command...> — gpurun (GPUS=N in the environment for --gpus N) with retries while the pod answers "busy / transient"
T=$1; LOG=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout "$T" -- "$@" > "$LOG" 2>&1
  if ! grep -q "status=transient\|rc=3\|no box\|busy" "$LOG"; then exit 0; fi
  sleep 90
done
