# the default bench line after attaching the ncu issue figures to the roofline block (bench.py changed: run it once as the driver does)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python bench.py > gpurun_out/r2_bench_final_tree_default.json 2> gpurun_out/r26_bench_err.txt; echo "rc=$?" >> gpurun_out/r26_bench_err.txt
tail -c 400 gpurun_out/r2_bench_final_tree_default.json
