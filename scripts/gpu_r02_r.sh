set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fused_train_gpu.py -q -k "sa2_fwd or sa2_bwd or sa2_wgrad" > gpurun_out/r02_r_k.log 2>&1; echo "kernel tests exit $?"; grep -E "passed|failed" gpurun_out/r02_r_k.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_r_k.log | cut -c1-260 | head -40
timeout 900 python -m pytest tests/test_fused_train_gpu.py -q -k "two_layer" > gpurun_out/r02_r_m.log 2>&1; echo "module tests exit $?"; grep -E "passed|failed" gpurun_out/r02_r_m.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_r_m.log | cut -c1-260 | head -30
