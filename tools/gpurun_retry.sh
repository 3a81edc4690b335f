#!/bin/bash
# gpurun_retry.sh <tag> <timeout_s> <command...>: retry while the pod answers busy (exit 3); log to gpurun_out/<tag>_call.log
TAG=$1; TMO=$2; shift 2
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun ${GPURUN_GPUS:+--gpus $GPURUN_GPUS} --timeout $TMO -- "$@" > gpurun_out/${TAG}_call.log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
