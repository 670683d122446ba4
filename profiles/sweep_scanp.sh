#!/bin/bash
# Tuning sweep of scanp_kernel: pairs of windows per trip (SPEQB_PWIN) x resident CTAs per SM (SPEQB_PMINB); rebuilds on the box.
O=${1:-gpurun_out/sweep}; mkdir -p $O
for v in ${SWEEP:-2:4 2:3 3:3 1:5 1:6}; do
  set -- ${v/:/ }
  SPEQB_NVCC_EXTRA="-DSPEQB_PWIN=$1 -DSPEQB_PMINB=$2" python -m speq_b200.build --force > /dev/null 2>&1
  for c in ${CONFIGS:-C2 C3}; do
    timeout 300 python bench.py --config $c --algorithm kpair --no-cpu --no-e2e $( [ $c = C4 ] && echo --reads 10000000 ) 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('pwin $1 minb $2 $c', round(d['roofline']['kernel_ms'],3), 'ms', round(d['value']/1e9,1), 'G/s', 'probes/window', round(d['roofline']['random_access']['sectors_per_launch']/d['tallies']['total_kmers'],3))" | tee -a $O/sweep_p.txt
  done
done
python -m speq_b200.build --force > /dev/null 2>&1
