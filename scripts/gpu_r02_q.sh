set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_geometry_gpu.py tests/test_fused_train_gpu.py -q > gpurun_out/r02_q_tests.log 2>&1; echo "tests exit $?"; grep -E "passed|failed" gpurun_out/r02_q_tests.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_q_tests.log | cut -c1-300 | head -20
timeout 600 python benchmarks/bench_ops.py --ops ball_query,three_interp --iters 5 --out gpurun_out/ops_bq_r02_q.jsonl > gpurun_out/ops_bq_r02_q.log 2>&1; echo "ops exit $?"; grep '"op": "ball_query"' gpurun_out/ops_bq_r02_q.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['N'], d['ms'], 'tiled', d['tiled_kernel_ms'], 'ref', d['ref_cuda_ms'], 'x', d['speedup_vs_ref_cuda'], d['bit_equal_to_tiled'])
"
