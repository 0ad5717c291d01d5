#!/bin/bash
# gpurun with retries while the pod answers "transient" (busy / draining): nothing is charged for those.
#   scripts/gpurun_retry.sh <timeout_s> '<command>' [gpus]
T=$1; CMD=$2; G=${3:-1}
for i in $(seq 1 40); do
  if [ "$G" = "1" ]; then OUT=$(/usr/local/graft/bin/gpurun --timeout $T -- "$CMD" 2>&1); else OUT=$(/usr/local/graft/bin/gpurun --gpus $G --timeout $T -- "$CMD" 2>&1); fi
  RC=$?
  echo "$OUT" | tail -40
  if echo "$OUT" | grep -q "status=transient"; then echo "[retry $i] transient, sleeping"; sleep 120; continue; fi
  if [ $RC -eq 3 ]; then echo "[retry $i] rc=3, sleeping"; sleep 120; continue; fi
  exit $RC
done
exit 3
