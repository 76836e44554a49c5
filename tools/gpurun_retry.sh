#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <timeout> [--gpus N] <command...>   (retries while the pod answers "transient"/busy)
LOG=$1; shift; TMO=$1; shift
G=""
if [ "$1" == "--gpus" ]; then G="--gpus $2"; shift; shift; fi
for i in 1 2 3 4 5 6 7 8 9 10; do
  /usr/local/graft/bin/gpurun $G --timeout $TMO -- "$@" > $LOG 2>&1
  if grep -q "status=transient\|exit code 3\|status=busy\|rc=3" $LOG; then sleep 90; continue; fi
  break
done
