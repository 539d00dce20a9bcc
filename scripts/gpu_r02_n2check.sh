set -x
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_r02_final2_n2.json 2> gpurun_out/bench_r02_final2_n2.err; echo "bench N=2 exit $?"; grep '^{' gpurun_out/bench_r02_final2_n2.json | cut -c1-260
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29532 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/bench_r02_final2_ref_n2.json 2> /dev/null; echo "ref N=2 exit $?"; grep '^{' gpurun_out/bench_r02_final2_ref_n2.json | cut -c1-200
