#!/bin/bash
# gpurun with retries on "transient" (exit 3: pod draining / no box): usage tools/gpurun_retry.sh [gpurun args...]
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
  /usr/local/graft/bin/gpurun "$@"; rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
