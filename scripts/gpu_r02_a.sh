# First GPU call of round 2 (one B200, ~3 min): parity tests, the contract bench, then the graph-captured model steps
# that round 1 left unmeasured (gap 6).  N>1: bash scripts/gpu_r01_an.sh N "b-cls xl-seg".
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -x -q -m gpu > gpurun_out/r02_a_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r02_a_tests.log
timeout 200 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r02_a.json 2> gpurun_out/bench_r02_a.err; echo "bench exit $?"; cut -c1-260 gpurun_out/bench_r02_a.json
for c in b-cls xl-seg; do
  timeout 150 python benchmarks/bench_models.py --steps 10 --warmup 3 --cases $c --graph >> gpurun_out/models_r02_a_graph_n1.jsonl 2> gpurun_out/models_r02_a_graph_$c.err; echo "$c graph exit $?"; tail -3 gpurun_out/models_r02_a_graph_$c.err
  timeout 150 python benchmarks/bench_models.py --steps 10 --warmup 3 --cases $c >> gpurun_out/models_r02_a_eager_n1.jsonl 2> gpurun_out/models_r02_a_eager_$c.err; echo "$c eager exit $?"
done
cut -c1-330 gpurun_out/models_r02_a_graph_n1.jsonl gpurun_out/models_r02_a_eager_n1.jsonl
