mkdir -p gpurun_out /tmp/ic; export SPEQ_BENCH_INDEX_CACHE=/tmp/ic
run() { # name, config, reads, env...
  n=$1; c=$2; r=$3; shift 3
  env "$@" timeout 600 python bench.py --config $c --reads $r --steps 3 --warmup 2 --no-cpu --no-e2e > gpurun_out/t_$n.json 2> gpurun_out/t_$n.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/t_$n.json')); print('$n', '%.4g'%d['value'], 'kern_ms', round(d['roofline']['kernel_ms'],3))"
}
run c4_base C4 10000000 X=1
run c4_t1 C4 10000000 SPEQ_TUNE=1
run c4_t2 C4 10000000 SPEQ_TUNE=2
run c4_g32 C4 10000000 SPEQ_L2_FETCH=32
run c4_g128 C4 10000000 SPEQ_L2_FETCH=128
run c4_g32t2 C4 10000000 SPEQ_L2_FETCH=32 SPEQ_TUNE=2
run c2_base C2 10000000 X=1
run c2_t1 C2 10000000 SPEQ_TUNE=1
run c2_t2 C2 10000000 SPEQ_TUNE=2
run c3_base C3 10000000 X=1
run c3_t1 C3 10000000 SPEQ_TUNE=1
run c3_g32 C3 10000000 SPEQ_L2_FETCH=32
