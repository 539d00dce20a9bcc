set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_j.json 2> gpurun_out/bench_r01_j.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_j.json
timeout 200 ncu --set full --clock-control none --import-source on -k regex:fps_kernel -c 1 -f -o gpurun_out/fps_2048_full_j python scripts/ncu_targets.py fps 2048 1024 > gpurun_out/ncu_fps.log 2>&1; echo "ncu fps exit $?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:group_rows -c 1 -f -o gpurun_out/group_16k_c128_full_j python scripts/ncu_targets.py group 16384 4096 128 > gpurun_out/ncu_group.log 2>&1; echo "ncu group exit $?"
timeout 900 python benchmarks/bench_ops.py --sizes 1024,2048,4096,8192,16384,32768,65536,100000 --out gpurun_out/ops_j.jsonl > /dev/null 2> gpurun_out/ops_j.err; echo "ops exit $?"; tail -3 gpurun_out/ops_j.err
python - <<'PY'
import json
for l in open('gpurun_out/ops_j.jsonl'):
    d=json.loads(l); print(d['op'].ljust(16), d['N'], d['ms'], d['frac_hbm'], d.get('speedup_vs_ref_cuda'), d.get('Tpairs_per_s_upper',''))
PY
