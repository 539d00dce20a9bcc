set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_p.json 2> gpurun_out/bench_r01_p.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_p.json; tail -3 gpurun_out/bench_r01_p.err
timeout 600 python benchmarks/bench_ops.py --no-ref --ops group_grad --sizes 1024,4096,16384,32768,65536,100000 --out gpurun_out/ops_p_grad.jsonl > /dev/null 2> gpurun_out/ops_p.err; echo "ops exit $?"; tail -3 gpurun_out/ops_p.err
python - <<'PY'
import json
for l in open('gpurun_out/ops_p_grad.jsonl'):
    d=json.loads(l); print(d['op'].ljust(16), d['N'], d['ms'], d['frac_hbm'], d.get('ms_l2_reductions'))
PY
timeout 200 ncu --set full --clock-control none --import-source on -k regex:group_grad_rows -c 1 -f -o gpurun_out/group_grad_16k_c32_full_p python scripts/ncu_targets.py group_grad 16384 4096 32 > gpurun_out/ncu_ggrad.log 2>&1; echo "ncu ggrad exit $?"
