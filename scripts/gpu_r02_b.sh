# Round 2, call b: the TMA-fed GEMM (tests + microbench), tc-profile goldens, graph tests, discriminator probe.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tcconv_gpu.py -q > gpurun_out/r02_b_tcconv.log 2>&1; echo "tcconv exit $?"; tail -15 gpurun_out/r02_b_tcconv.log
timeout 300 python scripts/bench_gemm.py > gpurun_out/gemm_r02_b.jsonl 2> gpurun_out/gemm_r02_b.err; echo "gemm bench exit $?"; cut -c1-150 gpurun_out/gemm_r02_b.jsonl; tail -3 gpurun_out/gemm_r02_b.err
timeout 900 python -m pytest tests/test_models_gpu.py tests/test_train_gpu.py tests/test_attention_gpu.py tests/test_norm_gpu.py -q -s > gpurun_out/r02_b_models.log 2>&1; echo "models exit $?"; grep -E "passed|failed" gpurun_out/r02_b_models.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_b_models.log | head -40
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_r02_b.json 2> gpurun_out/bench_r02_b.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r02_b.json; tail -3 gpurun_out/bench_r02_b.err
