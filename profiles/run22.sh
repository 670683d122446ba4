# strong scaling of the default bench on N GPUs of one box (N = $1), launched as the driver launches it
set -x
cd $GRAFT_REPO_ROOT
N=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench_f_C3_n$N.json 2> gpurun_out/r22_n${N}_err.txt
tail -c 600 gpurun_out/r2_bench_f_C3_n$N.json
