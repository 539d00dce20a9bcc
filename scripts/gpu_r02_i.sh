set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_tcconv_gpu.py tests/test_parity_gpu.py -q -k "conv or wgrad or interp" > gpurun_out/r02_i_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r02_i_tests.log | cut -c1-300
timeout 600 python scripts/bench_wgrad.py > gpurun_out/wgrad_r02_i.jsonl 2> gpurun_out/wgrad_r02_i.err; echo "wgrad bench exit $?"; tail -2 gpurun_out/wgrad_r02_i.err
