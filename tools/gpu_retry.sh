#!/usr/bin/env bash
# retry a gpurun call while the pod answers "busy" (exit code 3: nothing charged); usage: tools/gpu_retry.sh <timeout> <command>
T="$1"; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$T" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
