set -x
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_aa.json 2> gpurun_out/bench_r01_aa.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_aa.json; tail -3 gpurun_out/bench_r01_aa.err
