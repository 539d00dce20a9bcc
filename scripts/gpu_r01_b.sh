set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
nproc
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_b.json 2> gpurun_out/bench_r01_b.err; echo "bench exit $?"; tail -c 3000 gpurun_out/bench_r01_b.json; tail -5 gpurun_out/bench_r01_b.err
python benchmarks/bench_ops.py --out gpurun_out/ops_r01_b.jsonl > gpurun_out/ops_r01_b.log 2>&1; echo "ops exit $?"; tail -3 gpurun_out/ops_r01_b.log
