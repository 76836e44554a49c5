#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <timeout> <command...>   (retries while the pod answers "transient"/busy)
LOG=$1; shift; TMO=$1; shift
for i in 1 2 3 4 5 6 7 8; do
  /usr/local/graft/bin/gpurun --timeout $TMO -- "$@" > $LOG 2>&1
  if grep -q "status=transient\|exit code 3\|status=busy" $LOG; then sleep 60; continue; fi
  break
done
