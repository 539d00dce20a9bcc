set -x
mkdir -p gpurun_out
timeout 300 python scripts/dbg_tf32_round.py > gpurun_out/r02_c_dbg.log 2>&1; echo "dbg exit $?"; grep -v Warning gpurun_out/r02_c_dbg.log | tail -12
timeout 600 python -m pytest tests/test_fused_train_gpu.py tests/test_tcconv_gpu.py tests/test_models_gpu.py -q > gpurun_out/r02_c_tests.log 2>&1; echo "tests exit $?"; grep -E "passed|failed" gpurun_out/r02_c_tests.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_c_tests.log | cut -c1-300 | head -30
