# Last call of round 2: whole GPU suite, smoke, contract bench on the final tree.
set -x
mkdir -p gpurun_out
T=r02_final4
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/${T}_tests.log; grep -E "^FAILED|^E  " gpurun_out/${T}_tests.log | cut -c1-250 | head -10
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/${T}_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 --dump-kernels gpurun_out/kernels_${T}_n1.json > gpurun_out/bench_${T}_n1.json 2> gpurun_out/bench_${T}_n1.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_${T}_n1.json
timeout 600 python benchmarks/bench_models.py --steps 10 --warmup 3 > gpurun_out/models_${T}_n1.jsonl 2> gpurun_out/models_${T}_n1.err; echo "models exit $?"; cut -c1-200 gpurun_out/models_${T}_n1.jsonl
