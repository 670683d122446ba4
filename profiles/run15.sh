# walking scan, third build: GPU tests, ncu of scanw_kernel, A/B builds
set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -q -m gpu --maxfail=4 2>&1 | tail -40 > $O/r15_pytest.txt
tail -3 $O/r15_pytest.txt
F="--steps 2 --warmup 1 --no-cpu --no-e2e --no-probe"
cap() {   # name, kernel regex, skip, count, bench args...
  name=$1; rx=$2; skip=$3; cnt=$4; shift 4
  python bench.py "$@" $F > /dev/null 2>&1 || { echo "$name: plain run failed"; return; }
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -o /tmp/$name python bench.py "$@" $F > $O/ncu_$name.log 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > $O/${name}_raw.csv 2>/dev/null
  ncu -i /tmp/$name.ncu-rep --page source --csv > $O/${name}_source.csv 2>/dev/null
  ls -la /tmp/$name.ncu-rep $O/${name}_raw.csv $O/${name}_source.csv | awk '{print $5, $9}'
}
cap r2c_scanw_C3 scanw_kernel 4 2 --config C3 --reads 2000000
cap r2c_scanw_C4 scanw_kernel 4 2 --config C4 --reads 2000000
export CONFIGS="C2:10000000 C3:25000000 C4:20000000"
SWEEP="base;-DSPEQB_WMINB=2;-DSPEQB_KWIN=4;-DSPEQB_KWIN=4 -DSPEQB_WMINB=2" bash profiles/sweep_r2.sh $O
mv $O/r2_sweep.txt $O/r15_sweep.txt
SPEQB_FORCE_FILTER=1 SWEEP="base" bash profiles/sweep_r2.sh $O
mv $O/r2_sweep.txt $O/r15_sweep_filter.txt
du -sh $O
