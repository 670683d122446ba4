#!/bin/bash
# usage: gpu_retryN.sh <gpus> <outfile> <timeout> <command...>   (multi-GPU box; retries while the pod answers "busy")
n=$1; shift; out=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --gpus $n --timeout $to -- "$@" > $out 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (attempt $i)" >> $out; exit $rc; fi
  sleep 150
done
exit 3
