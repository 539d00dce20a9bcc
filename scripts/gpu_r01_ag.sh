set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 200 python bench.py --steps 10 --warmup 3 --dump-kernels gpurun_out/kernels_r01_ag.json > gpurun_out/bench_r01_ag.json 2> gpurun_out/bench_r01_ag.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_ag.json; tail -3 gpurun_out/bench_r01_ag.err
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
