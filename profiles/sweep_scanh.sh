#!/bin/bash
# Tuning sweep of scanh_kernel: windows per trip (SPEQB_KWIN) x resident CTAs per SM (SPEQB_HMINB); rebuilds the library on the box.
O=${1:-gpurun_out/sweep}; mkdir -p $O
for v in ${SWEEP:-3:4 2:4 4:3 2:5 4:2}; do
  set -- ${v/:/ }
  SPEQB_NVCC_EXTRA="-DSPEQB_KWIN=$1 -DSPEQB_HMINB=$2" python -m speq_b200.build --force > /dev/null 2>&1
  for c in C2 C3; do
    timeout 300 python bench.py --config $c --no-cpu --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('kwin $1 minb $2 $c', round(d['roofline']['kernel_ms'],3), 'ms', round(d['value']/1e9,1), 'G/s')" | tee -a $O/sweep.txt
  done
done
python -m speq_b200.build --force > /dev/null 2>&1
