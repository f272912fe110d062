#!/bin/bash
# tools/gpurun_retry.sh TIMEOUT 'command' -- gpurun with retries while the pod answers "no box free" (exit 3).
# GPUS=N in the environment asks for N GPUs of one box.
T=$1; shift
G=""; [ -n "$GPUS" ] && G="--gpus $GPUS"
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $G --timeout "$T" -- "$@"; rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
