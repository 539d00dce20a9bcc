set -x
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python bench.py --steps 10 --warmup 3 --dump-kernels gpurun_out/kernels_r01_af.json > gpurun_out/bench_r01_af.json 2> gpurun_out/bench_r01_af.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_af.json; tail -3 gpurun_out/bench_r01_af.err
python - <<'PY'
import json
for r in json.load(open('gpurun_out/kernels_r01_af.json')):
    if r['call']=='apb_group_grad_f32': print(r['dims'], r['launches'], r['us_per_launch'])
PY
timeout 300 python benchmarks/bench_ops.py --no-ref --ops group_grad --sizes 1024,16384,32768 --out gpurun_out/ops_af.jsonl > /dev/null 2> gpurun_out/ops_af.err; echo "ops exit $?"
python - <<'PY'
import json
for l in open('gpurun_out/ops_af.jsonl'):
    d=json.loads(l); print(d['op'].ljust(16), d['N'], d['ms'], d['frac_hbm'], d.get('ms_l2_reductions'))
PY
