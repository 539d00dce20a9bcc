set -x
mkdir -p gpurun_out
timeout 300 python scripts/dbg_wgrad.py > gpurun_out/r02_j_dbg.log 2>&1; cat gpurun_out/r02_j_dbg.log
timeout 300 python -m pytest tests/test_tcconv_gpu.py tests/test_parity_gpu.py -q -k "conv or wgrad or interp" > gpurun_out/r02_i_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r02_i_tests.log | cut -c1-300
timeout 600 python scripts/bench_wgrad.py > gpurun_out/wgrad_r02_j.jsonl 2> gpurun_out/wgrad_r02_j.err; echo "wgrad bench exit $?"; tail -2 gpurun_out/wgrad_r02_j.err
