#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> [gpurun args...] -- <command>
# Retries while the pod answers "busy/transient" (exit code 3); any other outcome ends the loop.
log=$1; shift
for attempt in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 60
done
exit 3
