set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_attention_gpu.py tests/test_parity_gpu.py tests/test_tcconv_gpu.py -q > gpurun_out/r02_h_tests.log 2>&1; echo "tests exit $?"; grep -E "passed|failed" gpurun_out/r02_h_tests.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_h_tests.log | cut -c1-300 | head -20
timeout 120 python scripts/bench_attn.py > gpurun_out/attn_r02_h.json 2>&1; cat gpurun_out/attn_r02_h.json
timeout 400 python bench.py --steps 20 --warmup 5 --no-cpu --dump-kernels gpurun_out/kernels_r02_h.json > gpurun_out/bench_r02_h.json 2> gpurun_out/bench_r02_h.err; echo "bench exit $?"; cut -c1-240 gpurun_out/bench_r02_h.json; tail -3 gpurun_out/bench_r02_h.err
timeout 900 python benchmarks/bench_ops.py --out gpurun_out/ops_microbench_r02_h.jsonl > gpurun_out/ops_r02_h.log 2>&1; echo "ops sweep exit $?"; tail -3 gpurun_out/ops_r02_h.log | cut -c1-300
