#!/bin/bash
# Retry wrapper around gpurun: exit code 3 (no box / slot free, nothing charged) is retried every 90 s.
#   scripts/grun.sh <log> [gpurun args...] -- '<command>'
log=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 90
done
exit 3
