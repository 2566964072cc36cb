#!/bin/bash
# gpurun with retries on "no slot right now" (exit 3): usage: tools/gpurun_retry.sh [gpurun args] -- 'command'
for attempt in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[gpurun_retry] attempt $attempt: no slot, sleeping 150 s" >&2
  sleep 150
done
exit 3
