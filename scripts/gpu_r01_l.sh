set -x
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_attention_gpu.py -x -q 2>&1 | tail -8
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_l.json 2> gpurun_out/bench_r01_l.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_l.json; tail -3 gpurun_out/bench_r01_l.err
timeout 200 python scripts/profile_step.py > gpurun_out/step_profile_l.txt 2>&1; head -9 gpurun_out/step_profile_l.txt
