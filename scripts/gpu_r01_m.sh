set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_m.json 2> gpurun_out/bench_r01_m.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_m.json; tail -3 gpurun_out/bench_r01_m.err
timeout 300 python bench.py --steps 10 --warmup 3 --no-fused-eval --no-cpu > gpurun_out/bench_r01_m_nofe.json 2> gpurun_out/bench_r01_m_nofe.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_m_nofe.json
timeout 200 python scripts/profile_step.py > gpurun_out/step_profile_m.txt 2>&1; head -9 gpurun_out/step_profile_m.txt
