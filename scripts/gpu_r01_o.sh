set -x
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_r01_o_n2.json 2> gpurun_out/bench_r01_o_n2.err; echo "n2 exit $?"; cut -c1-260 gpurun_out/bench_r01_o_n2.json; tail -5 gpurun_out/bench_r01_o_n2.err
timeout 600 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/bench_r01_o_ref.json 2> gpurun_out/bench_r01_o_ref.err; echo "ref exit $?"; cut -c1-300 gpurun_out/bench_r01_o_ref.json; tail -3 gpurun_out/bench_r01_o_ref.err
