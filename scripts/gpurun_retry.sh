#!/bin/bash
# usage: gpurun_retry.sh <gpus> <timeout> <logfile> <command...> — retries while the pod has no free slot (exit 3)
G=$1; T=$2; LOG=$3; shift 3
for i in $(seq 1 40); do
  if [ "$G" = "1" ]; then gpurun --timeout $T -- "$@" > $LOG 2>&1; else gpurun --gpus $G --timeout $T -- "$@" > $LOG 2>&1; fi
  rc=$?
  if grep -q "status=transient\|backing off\|exit code 3" $LOG || [ $rc -eq 3 ]; then sleep 120; continue; fi
  exit $rc
done
exit 3
