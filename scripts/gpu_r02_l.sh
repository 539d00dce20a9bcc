set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/r02_l_tests.log 2>&1; echo "tests exit $?"; grep -E "passed|failed" gpurun_out/r02_l_tests.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_l_tests.log | cut -c1-300 | head -20
timeout 400 python bench.py --steps 20 --warmup 5 --no-cpu --dump-kernels gpurun_out/kernels_r02_l.json > gpurun_out/bench_r02_l.json 2> gpurun_out/bench_r02_l.err; echo "bench exit $?"; cut -c1-240 gpurun_out/bench_r02_l.json; tail -3 gpurun_out/bench_r02_l.err
