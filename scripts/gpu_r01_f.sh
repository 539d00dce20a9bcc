set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -12
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_e.json 2> gpurun_out/bench_r01_e.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_r01_e.json; tail -5 gpurun_out/bench_r01_e.err
python scripts/profile_step.py > gpurun_out/step_profile_e.txt 2>&1; head -9 gpurun_out/step_profile_e.txt
