#!/bin/bash
# usage: gpu_retry.sh <outfile> <timeout> <command...>   (retries while the pod answers "busy", exit code 3)
out=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $out 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (attempt $i)" >> $out; exit $rc; fi
  sleep 150
done
exit 3
