# no-alloc entry point (alone in a fresh process, then after the other tests), C5 with the variant kept in walk-form slots,
# step-choice weights of the walking scan
set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu 2>&1 | tail -30 > $O/r18_pytest_noalloc_alone.txt
tail -3 $O/r18_pytest_noalloc_alone.txt
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -40 > $O/r18_pytest.txt
tail -3 $O/r18_pytest.txt
export CONFIGS="C2:10000000 C3:25000000 C4:20000000 C5:200000"
SWEEP="base;-DSPEQB_WGT_F=2 -DSPEQB_WGT_W=2;-DSPEQB_WGT_F=4 -DSPEQB_WGT_W=4;-DSPEQB_WGT_W=2;-DSPEQB_WGT_F=3 -DSPEQB_WGT_W=2;-DSPEQB_WMINB=4" bash profiles/sweep_r2.sh $O
mv $O/r2_sweep.txt $O/r18_sweep.txt
