# walking scan, second build: GPU parity tests, then A/B against the probing scans (SPEQB_NO_WALK=1)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -25 > gpurun_out/r14_pytest.txt
tail -3 gpurun_out/r14_pytest.txt
for nw in 0 1; do
  for c in C2:10000000 C3:25000000 C1:100000 C4:20000000; do
    set -- ${c/:/ }
    if [ $nw = 1 ]; then export SPEQB_NO_WALK=1; else unset SPEQB_NO_WALK; fi
    timeout 400 python bench.py --config $1 --reads $2 --cpu-seconds 3 --no-e2e --no-probe --steps 10 2>gpurun_out/r14_err_$1_$nw.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read()); t=d['tallies_only']
print('[no_walk=$nw] $1 record', round(d['roofline']['kernel_ms'],3), 'ms', round(d['value']/1e9,1), 'G/s | tallies only', round(t['kernel_ms'],3), 'ms', round(t['value']/1e9,1), 'G/s | probes', d['roofline']['sectors_per_launch'], '| parity', d.get('parity'))" | tee -a gpurun_out/r14_walk_ab.txt
  done
done
