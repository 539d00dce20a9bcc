# Round 2, 2-GPU call: bench.py must EXIT through the normal interpreter path at N>1 (no os._exit), peer all-reduce tests,
# cfg4 data-parallel graph step at N=2.
set -x
mkdir -p gpurun_out
N=${1:-2}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_r02_n$N.json 2> gpurun_out/bench_r02_n$N.err; echo "bench N=$N exit $?"; cut -c1-260 gpurun_out/bench_r02_n$N.json; grep -v "^W\|^\[W\|Warning" gpurun_out/bench_r02_n$N.err | tail -5
timeout 200 python -m pytest tests/test_peer_gpu.py -q > gpurun_out/r02_n${N}_peer.log 2>&1; echo "peer tests exit $?"; tail -3 gpurun_out/r02_n${N}_peer.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 benchmarks/bench_models.py --gpus $N --steps 10 --warmup 3 --cases b-cls,xl-seg > gpurun_out/models_r02_n$N.jsonl 2> gpurun_out/models_r02_n$N.err; echo "models N=$N exit $?"; cut -c1-200 gpurun_out/models_r02_n$N.jsonl; grep -v "^W\|^\[W\|Warning" gpurun_out/models_r02_n$N.err | tail -5
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_r02_ref_n$N.json 2> /dev/null; echo "ref arm N=$N exit $?"; cut -c1-200 gpurun_out/bench_r02_ref_n$N.json
