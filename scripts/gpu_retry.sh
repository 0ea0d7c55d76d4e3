#!/bin/bash
# usage: gpu_retry.sh LOGFILE TIMEOUT 'command'   -- retries while the pod answers busy (exit 3 / transient)
LOG=$1; TMO=$2; CMD=$3
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $TMO -- "$CMD" > $LOG 2>&1
  if ! grep -q "status=transient\|rc=None" $LOG; then break; fi
  sleep 45
done
echo "gpu_retry done" >> $LOG
