#!/bin/bash
# Round-2 ncu captures on the GPU box (one GPU).  Every capture runs the same command without ncu first; the reports are
# exported to CSV (raw + source pages) and deleted, so that gpurun_out/ stays below the 64 MiB that is copied back.
O=gpurun_out; mkdir -p $O
F="--steps 2 --warmup 1 --no-cpu --no-e2e --no-probe"
cap() {   # name, kernel regex, skip, count, bench args...
  name=$1; rx=$2; skip=$3; cnt=$4; shift 4
  python bench.py "$@" $F > /dev/null 2>&1 || { echo "$name: plain run failed"; return; }
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -o /tmp/$name python bench.py "$@" $F > $O/ncu_$name.log 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > $O/${name}_raw.csv 2>/dev/null
  ncu -i /tmp/$name.ncu-rep --page source --csv > $O/${name}_source.csv 2>/dev/null
  ls -la /tmp/$name.ncu-rep $O/${name}_raw.csv $O/${name}_source.csv | awk '{print $5, $9}'
}
B="python bench.py --config C3 --reads 20000000 --steps 2 --warmup 1 --no-cpu"
# launch list of the library's own kernels in one bench run (torch's generator kernels of the synthetic workload are filtered out)
$B > $O/r2_plain_C3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"scan|pack|fq_|kpos|ktable|kpair|kfilter|fill|refine|probe|build_up|sa_|count_nonzero" -c 3000 --csv --log-file $O/r2_launches_C3.csv $B > $O/r2_ncu_launches.log 2>&1
cap r2_scanh_C3 scanh_kernel 4 2 --config C3 --reads 2000000
cap r2_scanh_C2 scanh_kernel 4 2 --config C2 --reads 1000000
cap r2_scan32_local_k70 scan32_kernel 2 1 --config C2 --reads 1000000 --mode local -k 70
cap r2_scanh_local scanh_kernel 2 1 --config C2 --reads 1000000 --mode local
cap r2_scanp_C4 scanp_kernel 4 2 --config C4 --reads 2000000
du -sh $O
