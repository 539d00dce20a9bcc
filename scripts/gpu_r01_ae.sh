set -x
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python bench.py --steps 10 --warmup 3 --dump-kernels gpurun_out/kernels_r01_ae.json > gpurun_out/bench_r01_ae.json 2> gpurun_out/bench_r01_ae.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_ae.json; tail -3 gpurun_out/bench_r01_ae.err
python - <<'PY'
import json
for r in json.load(open('gpurun_out/kernels_r01_ae.json')):
    if r['call']=='apb_group_grad_f32': print(r['dims'], r['launches'], r['us_per_launch'])
PY
