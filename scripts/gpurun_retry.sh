#!/bin/bash
# gpurun with retry on "no box / slot free" (exit 3) and "another call running" (exit 2 is NOT retried).
# usage: scripts/gpurun_retry.sh [--gpus N] [--timeout S] -- 'command'
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
