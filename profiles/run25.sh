# last check of the final tree: all GPU tests, smoke(), the default bench line
set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 > $O/r25_pytest.txt
tail -2 $O/r25_pytest.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r25_smoke.txt 2>&1
python bench.py > $O/r2_bench_final_tree_default.json 2> $O/r25_bench_err.txt
tail -c 300 $O/r2_bench_final_tree_default.json
