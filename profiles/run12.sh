set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_cli.py -q -m gpu -x 2>&1 | tail -15 > gpurun_out/r12_pytest_cli.txt
timeout 600 python profiles/cli_e2e.py 10000000 paired > gpurun_out/r2_cli_e2e_C2_10M_paired.json 2> gpurun_out/r12_cli_paired.err
tail -3 gpurun_out/r12_pytest_cli.txt
