#!/bin/bash
# usage: scripts/gpu_retry.sh <timeout> <out-log> [--gpus N] <command...> ; retries while the pod answers "busy" (exit 3)
T=$1; LOG=$2; shift 2
G=""
if [ "$1" == "--gpus" ]; then G="--gpus $2"; shift 2; fi
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $G --timeout $T -- "$@" > $LOG 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" $LOG; then exit $rc; fi
  sleep 120
done
exit 3
