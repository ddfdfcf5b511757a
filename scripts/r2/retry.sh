#!/bin/bash
# usage: scripts/r2/retry.sh <timeout> [--gpus N] -- <command>   (re-submits while gpurun answers "busy", rc 3)
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[retry] attempt $attempt answered busy; sleeping 120 s"
  sleep 120
done
exit 3
