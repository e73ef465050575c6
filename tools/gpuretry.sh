#!/bin/bash
# usage: tools/gpuretry.sh <logfile> <timeout> <cmd...>   -- retries gpurun while the pod answers busy (exit 3)
LOG=$1; TO=$2; shift 2
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $TO -- "$@" > $LOG 2>&1
  rc=$?
  echo "[gpuretry] attempt $i rc=$rc $(date)" >> $LOG.attempts
  if [ $rc -ne 3 ]; then echo "[gpuretry] done rc=$rc" >> $LOG; exit $rc; fi
  sleep 90
done
