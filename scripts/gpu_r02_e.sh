# Round 2, call e: training-mode fused one-conv set abstraction (tests), goldens on tc, cfg4/cfg5 graph steps at SURVEY 8d batches
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fused_train_gpu.py tests/test_tcconv_gpu.py tests/test_models_gpu.py tests/test_train_gpu.py -q > gpurun_out/r02_e_tests.log 2>&1; echo "tests exit $?"; grep -E "passed|failed" gpurun_out/r02_e_tests.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_e_tests.log | head -30
for c in b-cls l-cls xl-seg; do
  timeout 300 python benchmarks/bench_models.py --steps 10 --warmup 3 --cases $c >> gpurun_out/models_r02_e_n1.jsonl 2> gpurun_out/models_r02_e_$c.err; echo "$c exit $?"; tail -2 gpurun_out/models_r02_e_$c.err
done
cut -c1-420 gpurun_out/models_r02_e_n1.jsonl
APB_NO_FUSED_TRAIN=1 timeout 300 python benchmarks/bench_models.py --steps 10 --warmup 3 --cases b-cls,xl-seg > gpurun_out/models_r02_e_n1_unfused.jsonl 2> gpurun_out/models_r02_e_unfused.err; echo "unfused exit $?"; cut -c1-300 gpurun_out/models_r02_e_n1_unfused.jsonl; tail -2 gpurun_out/models_r02_e_unfused.err
