set -x
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_r01_t_n2.json 2> gpurun_out/bench_r01_t_n2.err; echo "n2 exit $?"; cut -c1-260 gpurun_out/bench_r01_t_n2.json; tail -5 gpurun_out/bench_r01_t_n2.err
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/bench_r01_t_ref_n2.json 2> gpurun_out/bench_r01_t_ref_n2.err; echo "ref n2 exit $?"; cut -c1-200 gpurun_out/bench_r01_t_ref_n2.json
