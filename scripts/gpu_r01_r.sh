set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 600 python benchmarks/bench_ops.py --no-ref --ops group,group_grad --sizes 65536,100000 --out gpurun_out/ops_r.jsonl > /dev/null 2> gpurun_out/ops_r.err; echo "ops exit $?"; tail -3 gpurun_out/ops_r.err
python - <<'PY'
import json
for l in open('gpurun_out/ops_r.jsonl'):
    d=json.loads(l); print(d['op'].ljust(16), d['N'], d['ms'], d['frac_hbm'], d.get('ms_l2_reductions'), d.get('ms_direct_l2_gather'))
PY
timeout 200 ncu --set full --clock-control none --import-source on -k regex:group_rows_seg -c 1 -f -o gpurun_out/group_100k_c32_seg_r python scripts/ncu_targets.py group 100000 25000 32 > gpurun_out/ncu_gseg.log 2>&1; echo "ncu seg exit $?"
