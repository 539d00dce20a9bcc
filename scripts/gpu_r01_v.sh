set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_r01_v_n8.json 2> gpurun_out/bench_r01_v_n8.err; echo "n8 exit $?"; cut -c1-260 gpurun_out/bench_r01_v_n8.json; tail -5 gpurun_out/bench_r01_v_n8.err
