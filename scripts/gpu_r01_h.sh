set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q 2>&1 | tail -4
python benchmarks/bench_ops.py --sizes 4096,16384,32768 --ops group,gather --tune --no-ref --out gpurun_out/ops_h_tune.jsonl > /dev/null 2> gpurun_out/ops_h.err; echo "ops exit $?"; tail -3 gpurun_out/ops_h.err
python - <<'PY'
import json
for l in open('gpurun_out/ops_h_tune.jsonl'):
    d=json.loads(l); print(d['op'].ljust(28), d['N'], d['ms'], d['frac_hbm'])
PY
