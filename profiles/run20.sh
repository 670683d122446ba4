# cooperative probe trips of the walking scan: GPU tests, then A/B builds
set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -30 > $O/r20_pytest.txt
tail -3 $O/r20_pytest.txt
export CONFIGS="C2:10000000 C3:25000000 C4:20000000 C1:100000"
SWEEP="base;-DSPEQB_WCOOP=0;-DSPEQB_WMINB=4;-DSPEQB_WGT_F=1 -DSPEQB_WGT_W=1;-DSPEQB_WGT_F=2 -DSPEQB_WGT_W=1" bash profiles/sweep_r2.sh $O
mv $O/r2_sweep.txt $O/r20_sweep.txt
