#!/bin/bash
# usage: tools/gpu_retry.sh <logfile> <timeout-seconds> <command...>   - retries while the pod answers "busy" (exit code 3)
LOG=$1; shift; TMO=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout $TMO -- "$@" > $LOG 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "rc=$rc" >> $LOG; exit $rc; fi
  sleep 90
done
echo "rc=3 (gave up)" >> $LOG
