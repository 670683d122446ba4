#!/bin/bash
# Round-2 A/B builds of the table scan on the GPU box: each variant rebuilds the library with SPEQB_NVCC_EXTRA and prints the
# record-mode (value) and tallies-only kernel times of `bench.py --config C2` and C3 at 25 M reads.
O=${1:-gpurun_out}; mkdir -p $O
IFS=";" read -ra VS <<< "${SWEEP:-base}"
for v in "${VS[@]}"; do
  extra=""; [ "$v" != "base" ] && extra="$v"
  SPEQB_NVCC_EXTRA="$extra" python -m speq_b200.build --force > /dev/null 2>&1 || { echo "build failed: $v" | tee -a $O/r2_sweep.txt; continue; }
  for c in ${CONFIGS:-C2:10000000 C3:25000000}; do
    set -- ${c/:/ }
    timeout 300 python bench.py --config $1 --reads $2 --no-cpu --no-e2e --no-probe --steps 10 $BENCH_ARGS 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); t=d['tallies_only']
print('[$v] $BENCH_ARGS $1 record', round(d['roofline']['kernel_ms'],3), 'ms', round(d['value']/1e9,1), 'G/s | tallies only', round(t['kernel_ms'],3), 'ms', round(t['value']/1e9,1), 'G/s')" | tee -a $O/r2_sweep.txt
  done
done
python -m speq_b200.build --force > /dev/null 2>&1
