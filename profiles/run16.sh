# walking scan with one step per loop trip (the state most lanes wait in) and first-hit consume: tests, A/B builds, ncu
set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -q -m gpu --maxfail=4 2>&1 | tail -40 > $O/r16_pytest.txt
tail -3 $O/r16_pytest.txt
export CONFIGS="C2:10000000 C3:25000000 C4:20000000 C1:100000"
SWEEP="base;-DSPEQB_WKWIN=4;-DSPEQB_WKWIN=2;-DSPEQB_WKWIN=4 -DSPEQB_WMINB=2" bash profiles/sweep_r2.sh $O
mv $O/r2_sweep.txt $O/r16_sweep.txt
F="--steps 2 --warmup 1 --no-cpu --no-e2e --no-probe"
cap() {   # name, kernel regex, skip, count, bench args...
  name=$1; rx=$2; skip=$3; cnt=$4; shift 4
  python bench.py "$@" $F > /dev/null 2>&1 || { echo "$name: plain run failed"; return; }
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -o /tmp/$name python bench.py "$@" $F > $O/ncu_$name.log 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > $O/${name}_raw.csv 2>/dev/null
  ncu -i /tmp/$name.ncu-rep --page source --csv > $O/${name}_source.csv 2>/dev/null
  ls -la /tmp/$name.ncu-rep $O/${name}_raw.csv $O/${name}_source.csv | awk '{print $5, $9}'
}
cap r2d_scanw_C3 scanw_kernel 4 2 --config C3 --reads 2000000
du -sh $O
