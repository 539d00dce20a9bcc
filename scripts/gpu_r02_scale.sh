# Round 2 scaling call: bash scripts/gpu_r02_scale.sh N   (run under `gpurun --gpus N`)
# contract bench (cfg2) + its per-kernel / per-collective trace, 8-GPU peer all-reduce test, cfg4 / cfg5 model steps.
set -x
mkdir -p gpurun_out
N=${1:-8}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $RUN --master-port 29521 bench.py --gpus $N --steps 20 --warmup 5 --dump-kernels gpurun_out/kernels_r02_final5_n$N.json > gpurun_out/bench_r02_final5_n$N.json 2> gpurun_out/bench_r02_final5_n$N.err; echo "bench N=$N exit $?"; cut -c1-260 gpurun_out/bench_r02_final5_n$N.json; grep -v "^W\|^\[W\|Warning" gpurun_out/bench_r02_final5_n$N.err | tail -4
timeout 200 python -m pytest tests/test_peer_gpu.py -q -s > gpurun_out/r02_n${N}_peer.log 2>&1; echo "peer tests exit $?"; tail -3 gpurun_out/r02_n${N}_peer.log
timeout 400 $RUN --master-port 29522 benchmarks/bench_models.py --gpus $N --steps 10 --warmup 3 --cases b-cls,l-cls,xl-seg > gpurun_out/models_r02_final5_n$N.jsonl 2> gpurun_out/models_r02_final5_n$N.err; echo "models N=$N exit $?"; cut -c1-200 gpurun_out/models_r02_final5_n$N.jsonl; grep -v "^W\|^\[W\|Warning" gpurun_out/models_r02_final5_n$N.err | tail -4
