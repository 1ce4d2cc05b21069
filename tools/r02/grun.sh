#!/bin/bash
# usage: grun.sh <tag> <timeout_s> <command...>   -- gpurun with retries while the pod answers "busy" (exit 3)
tag=$1; shift; to=$1; shift
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > gpurun_out/call_$tag.log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "rc=$rc attempt=$attempt" >> gpurun_out/call_$tag.log; exit $rc; fi
  sleep 60
done
