set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 200 python bench.py --steps 10 --warmup 3 --dump-kernels gpurun_out/kernels_r01_ak.json > gpurun_out/bench_r01_ak.json 2> gpurun_out/bench_r01_ak.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_ak.json; tail -3 gpurun_out/bench_r01_ak.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_r01_ak.json'))
for f in d['our_kernels']['families'][:10]: print(f)
PY
