# checking build of the walking scan (-DSPEQB_BOUNDS=1): the global-mode GPU tests must pass without a failed index check
set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
SPEQB_NVCC_EXTRA="-DSPEQB_BOUNDS=1" python -m speq_b200.build --force > /dev/null 2>&1 || echo "build failed" > $O/r24_pytest_bounds.txt
timeout 1200 python -m pytest -q -m gpu -x tests/test_gpu_parity.py tests/test_gpu_oracle_shapes.py tests/test_cli.py 2>&1 | tail -8 >> $O/r24_pytest_bounds.txt
tail -3 $O/r24_pytest_bounds.txt
python bench.py --config C4 --reads 20000000 --steps 3 --warmup 1 --no-e2e --no-probe --cpu-seconds 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('C4 checking build: parity', d['parity']['ok'], 'value %.3e' % d['value'])" >> $O/r24_pytest_bounds.txt
python -m speq_b200.build --force > /dev/null 2>&1
