set -x
mkdir -p gpurun_out
N=${1:-1}
for c in ${2:-b-cls xl-seg l-cls}; do
  if [ "$N" = 1 ]; then
    timeout 150 python benchmarks/bench_models.py --gpus 1 --steps 5 --warmup 3 --cases $c >> gpurun_out/models_r01_an_n$N.jsonl 2> gpurun_out/models_r01_an_n${N}_$c.err
  else
    timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 benchmarks/bench_models.py --gpus $N --steps 5 --warmup 3 --cases $c >> gpurun_out/models_r01_an_n$N.jsonl 2> gpurun_out/models_r01_an_n${N}_$c.err
  fi
  echo "$c n$N exit $?"; tail -3 gpurun_out/models_r01_an_n${N}_$c.err
done
cut -c1-330 gpurun_out/models_r01_an_n$N.jsonl
if [ "$N" = 1 ]; then
  timeout 150 python benchmarks/bench_models.py --gpus 1 --steps 5 --warmup 3 --cases b-cls,xl-seg --conv torch > gpurun_out/models_r01_an_n1_cudnn.jsonl 2> gpurun_out/models_r01_an_n1_cudnn.err; echo "cudnn exit $?"
  cut -c1-330 gpurun_out/models_r01_an_n1_cudnn.jsonl; tail -3 gpurun_out/models_r01_an_n1_cudnn.err
fi
