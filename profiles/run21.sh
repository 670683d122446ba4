# final build f (walking scan, weights 3/2/1, own launch configuration): GPU tests, smoke, default bench, parity shapes, ncu
set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -q -m gpu --maxfail=4 2>&1 | tail -40 > $O/r21_pytest.txt
tail -3 $O/r21_pytest.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r21_smoke.txt 2>&1
t0=$(date +%s); python bench.py --steps 20 --warmup 5 > $O/r2_bench_f_C3_default.json 2> $O/r21_bench_err.txt; echo "wall $(( $(date +%s) - t0 )) s" >> $O/r21_bench_err.txt
F="--steps 2 --warmup 1 --no-cpu --no-e2e --no-probe"
cap() {   # name, kernel regex, skip, count, bench args...
  name=$1; rx=$2; skip=$3; cnt=$4; shift 4
  python bench.py "$@" $F > /dev/null 2>&1 || { echo "$name: plain run failed"; return; }
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -o /tmp/$name python bench.py "$@" $F > $O/ncu_$name.log 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > $O/${name}_raw.csv 2>/dev/null
  ncu -i /tmp/$name.ncu-rep --page source --csv > $O/${name}_source.csv 2>/dev/null
}
cap r2f_scanw_C3 scanw_kernel 4 2 --config C3 --reads 2000000
cap r2f_scanw_C4 scanw_kernel 4 2 --config C4 --reads 2000000
B="python bench.py --config C3 --reads 20000000 --steps 2 --warmup 1 --no-cpu"
$B > $O/r21_plain_C3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"scan|pack|fq_|kpos|ktable|kpair|kfilter|fill|refine|probe|build_up|sa_|count_nonzero" -c 3000 --csv --log-file $O/r2f_launches_C3.csv $B > $O/r21_ncu_launches.log 2>&1
du -sh $O
