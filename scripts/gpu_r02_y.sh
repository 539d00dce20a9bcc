set -x
mkdir -p gpurun_out
for d in 0 1 2 4 7; do
echo "dbg $d"; timeout 200 python scripts/bench_sa2.py --only bwd --dbg $d --iters 5 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l)
        if d['stage'] in (0, 3): print(d['stage'], d['training'], d['us'])
"
done
