set -x
mkdir -p gpurun_out
timeout 180 python tests/test_peer_gpu.py 2 > gpurun_out/peer_test_u.log 2>&1; echo "peer test exit $?"; tail -5 gpurun_out/peer_test_u.log
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_r01_u_n2_peer.json 2> gpurun_out/bench_r01_u_n2_peer.err; echo "n2 peer exit $?"; cut -c1-260 gpurun_out/bench_r01_u_n2_peer.json; tail -5 gpurun_out/bench_r01_u_n2_peer.err
APB_PEER_ALLREDUCE=0 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_r01_u_n2_nccl.json 2> gpurun_out/bench_r01_u_n2_nccl.err; echo "n2 nccl exit $?"; cut -c1-260 gpurun_out/bench_r01_u_n2_nccl.json; tail -5 gpurun_out/bench_r01_u_n2_nccl.err
