#!/bin/bash
# gpurun with retries while the pod answers "busy" (exit code 3) or an earlier call of ours is still registered:
#   tools/gpu_retry.sh [gpurun options] -- '<command>'
for i in $(seq 1 60); do
  while /usr/local/graft/bin/gpurun --status | grep -q '"in_flight": 1'; do sleep 10; done
  /usr/local/graft/bin/gpurun "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 30
done
exit 3
