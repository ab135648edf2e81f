#!/bin/bash
# gpurun with retries while the pod answers "busy" (exit 3, nothing charged).  usage: tools/gpu_retry.sh <log> <gpurun args...>
LOG=$1; shift
for i in $(seq 1 ${RETRIES:-40}); do
    /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
    rc=$?
    if [ $rc -ne 3 ]; then exit $rc; fi
    sleep ${RETRY_SLEEP:-45}
done
exit 3
