set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/r02_v_alone.log 2>&1; echo "alone exit $?"; tail -3 gpurun_out/r02_v_alone.log; grep -E "^E  " gpurun_out/r02_v_alone.log | cut -c1-250 | head
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02_v_all.log 2>&1; echo "all exit $?"; tail -3 gpurun_out/r02_v_all.log; grep -E "^FAILED|^E  " gpurun_out/r02_v_all.log | cut -c1-250 | head -20
