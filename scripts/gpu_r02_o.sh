set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/r02_o_tests.log 2>&1; echo "tests exit $?"; grep -E "passed|failed" gpurun_out/r02_o_tests.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_o_tests.log | cut -c1-300 | head -20
timeout 300 python benchmarks/bench_models.py --steps 10 --warmup 3 --cases b-cls,l-cls,xl-seg > gpurun_out/models_r02_o_n1.jsonl 2> gpurun_out/models_r02_o.err; echo "models exit $?"; cut -c1-260 gpurun_out/models_r02_o_n1.jsonl; tail -2 gpurun_out/models_r02_o.err
timeout 300 python benchmarks/bench_models.py --steps 10 --warmup 3 --cases b-cls,xl-seg --no-prefetch > gpurun_out/models_r02_o_n1_inline.jsonl 2> gpurun_out/models_r02_o2.err; echo "inline exit $?"; cut -c1-260 gpurun_out/models_r02_o_n1_inline.jsonl; tail -2 gpurun_out/models_r02_o2.err
