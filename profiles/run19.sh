# build e (walking scan with its own launch configuration, weights 3/2/1): GPU tests, bench lines of all shapes, C5 A/B, CLI timing
set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -30 > $O/r19_pytest.txt
tail -3 $O/r19_pytest.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r19_smoke.txt 2>&1
t0=$(date +%s); python bench.py --steps 20 --warmup 5 > $O/r2_bench_e_C3_default.json 2> $O/r19_bench_err.txt; echo "wall $(( $(date +%s) - t0 )) s" >> $O/r19_bench_err.txt
for c in C2:10000000 C4:100000000 C5:1000000 C1:100000; do
  set -- ${c/:/ }
  timeout 600 python bench.py --config $1 --reads $2 --steps 10 --warmup 3 > $O/r2_bench_e_$1.json 2> $O/r19_err_$1.txt
done
SPEQB_NO_WALK=1 timeout 600 python bench.py --config C5 --reads 1000000 --steps 10 --warmup 3 --no-e2e --no-probe --cpu-seconds 3 > $O/r19_C5_nowalk.json 2> $O/r19_err_C5_nowalk.txt
timeout 600 python profiles/cli_e2e.py 10000000 > $O/r2_cli_e2e_C2_10M_e.json 2> $O/r19_cli.err
timeout 600 python profiles/cli_e2e.py 10000000 paired > $O/r2_cli_e2e_C2_10M_paired_e.json 2> $O/r19_cli_paired.err
du -sh $O
