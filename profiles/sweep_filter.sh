#!/bin/bash
# Tuning sweep of the presence filter of the pair-table scan on C4: L2 eviction hints on/off x filter size (2^FBITS_MAX bits);
# rebuilds on the box.
O=${1:-gpurun_out/sweep_filter}; mkdir -p $O
for v in ${SWEEP:-1:28 1:27 1:26 0:27 1:29}; do
  set -- ${v/:/ }
  SPEQB_NVCC_EXTRA="-DSPEQB_L2HINT=$1 -DSPEQB_FBITS_MAX=$2" python -m speq_b200.build --force > /dev/null 2>&1
  timeout 300 python bench.py --config C4 --reads 10000000 --no-cpu --no-e2e > $O/l.json 2>/dev/null
  echo "hint $1 fbits $2: $(python profiles/bench_line.py $O/l.json)" | tee -a $O/sweep_filter.txt
done
python -m speq_b200.build --force > /dev/null 2>&1
