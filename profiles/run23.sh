# (compute-sanitizer turned out to be closed on this pool: the memcheck step below printed the pool's notice and ran nothing.)
# compute-sanitizer memcheck over the walking scan's differential test, the walk-form table builds and the no-alloc entry point;
# CLI tests and paired CLI timing after pinning the second staging array at start-up
set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 --log-file $O/r23_memcheck.txt python -m pytest -q -m gpu -x \
  "tests/test_gpu_parity.py::test_walking_scan_matches_rolling" "tests/test_gpu_parity.py::test_golden_ktable_algorithm" \
  "tests/test_gpu_parity.py::test_edge_cases" tests/test_zz_no_alloc.py > $O/r23_memcheck_pytest.txt 2>&1
echo "memcheck rc=$?" >> $O/r23_memcheck_pytest.txt
tail -5 $O/r23_memcheck_pytest.txt; tail -5 $O/r23_memcheck.txt
timeout 600 python -m pytest tests/test_cli.py -q -m gpu -x 2>&1 | tail -5 > $O/r23_pytest_cli.txt
tail -2 $O/r23_pytest_cli.txt
timeout 600 python profiles/cli_e2e.py 10000000 paired > $O/r2_cli_e2e_C2_10M_paired_f.json 2> $O/r23_cli_paired.err
timeout 600 python profiles/cli_e2e.py 10000000 > $O/r2_cli_e2e_C2_10M_f.json 2> $O/r23_cli.err
