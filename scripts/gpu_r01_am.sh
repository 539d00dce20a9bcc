set -x
mkdir -p gpurun_out
N=${1:-2}
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_r01_am_n$N.json 2> gpurun_out/bench_r01_am_n$N.err; echo "n$N exit $?"; cut -c1-260 gpurun_out/bench_r01_am_n$N.json; tail -4 gpurun_out/bench_r01_am_n$N.err
