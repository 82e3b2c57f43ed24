#!/bin/bash
# tests/gr.sh <timeout-seconds> <command> -- gpurun with retries while the pod is busy (exit code 3).  Tooling only.
T=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $T -- "$@" > /tmp/gr_last.log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then tail -40 /tmp/gr_last.log; exit $rc; fi
  sleep 60
done
echo "gave up"; exit 3
