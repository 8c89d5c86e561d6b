#!/bin/bash
# usage: tools/gpurun_retry.sh <log> <timeout-seconds> <gpus> <command...>
# retries a gpurun call while the pod answers "busy / transient" (nothing is charged for those)
LOG=$1; TMO=$2; GPUS=$3; shift 3
for i in $(seq 1 40); do
  if [ "$GPUS" = "1" ]; then gpurun --timeout $TMO -- "$@" > $LOG 2>&1; else gpurun --gpus $GPUS --timeout $TMO -- "$@" > $LOG 2>&1; fi
  if grep -q "status=transient\|status=busy\|exit code 3\|rc=3" $LOG; then sleep 120; else break; fi
done
tail -80 $LOG
