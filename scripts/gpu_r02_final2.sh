set -x
mkdir -p gpurun_out
T=r02_final2
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/${T}_tests.log; grep -E "^FAILED|^E  " gpurun_out/${T}_tests.log | cut -c1-250 | head -10
timeout 600 python bench.py --steps 20 --warmup 5 --dump-kernels gpurun_out/kernels_${T}_n1.json > gpurun_out/bench_${T}_n1.json 2> gpurun_out/bench_${T}_n1.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_${T}_n1.json
timeout 300 python scripts/profile_torch_ops.py > gpurun_out/torch_ops_${T}.txt 2> /dev/null; echo "profile exit $?"; head -6 gpurun_out/torch_ops_${T}.txt | cut -c1-200
