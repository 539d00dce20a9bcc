set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/tmp_k.log 2>&1; echo "tests exit $?"; tail -1 gpurun_out/tmp_k.log; grep -E "^FAILED|^ERROR|^E  " gpurun_out/tmp_k.log | cut -c1-260 | head -20
timeout 200 python bench.py --no-cpu --steps 20 --warmup 5 --dump-kernels gpurun_out/kernels_tmp.json > gpurun_out/bench_tmp.json 2> gpurun_out/bench_tmp.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_tmp.json; tail -3 gpurun_out/bench_tmp.err
timeout 400 python benchmarks/bench_models.py --steps 10 --warmup 3 > gpurun_out/models_tmp.jsonl 2> /dev/null; echo "models exit $?"; cut -c1-150 gpurun_out/models_tmp.jsonl
