set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fused_train_gpu.py -q -k "sa2_fwd or sa2_bwd or sa2_wgrad or two_layer" > gpurun_out/r02_za_k.log 2>&1; echo "kernel tests exit $?"; grep -E "passed|failed" gpurun_out/r02_za_k.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_za_k.log | cut -c1-260 | head -40
timeout 300 python scripts/bench_sa2.py --out gpurun_out/sa2_microbench_r02_za.jsonl > gpurun_out/sa2_z.log 2>&1; echo "microbench exit $?"; python - <<'PY'
import json
for l in open('gpurun_out/sa2_microbench_r02_za.jsonl'):
    d = json.loads(l); print(d['stage'], d['kernel'], d.get('training', d.get('gram', '')), d['us'], d['tflops'])
PY
for d in 7; do
echo "dbg $d"; timeout 200 python scripts/bench_sa2.py --only bwd --dbg $d --iters 5 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l)
        if d['stage'] in (0, 3): print('D', d['stage'], d['training'], d['us'])
"
done
timeout 200 python bench.py --no-cpu --dump-kernels gpurun_out/kernels_r02_za_fused2.json > gpurun_out/bench_r02_za_fused2.json 2> gpurun_out/bench_r02_za_fused2.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_r02_za_fused2.json; tail -3 gpurun_out/bench_r02_za_fused2.err
