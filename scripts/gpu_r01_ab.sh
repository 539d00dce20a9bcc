set -x
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_ac.json 2> gpurun_out/bench_r01_ac.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_ac.json; tail -3 gpurun_out/bench_r01_ac.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_r01_ac.json'))
for f in d['our_kernels']['families'][:6]: print(f)
PY
