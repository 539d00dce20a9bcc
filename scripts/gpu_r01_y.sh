set -x
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_y.json 2> gpurun_out/bench_r01_y.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_y.json; tail -3 gpurun_out/bench_r01_y.err
python bench.py --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/launches_r01_y.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_y.log 2>&1; echo "ncu list exit $?"
for t in "pwtf32" "pw" "sa2"; do
  python scripts/ncu_targets.py $t > /dev/null 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:tc_rows_gemm -c 1 -f -o gpurun_out/tc_${t}_full_y python scripts/ncu_targets.py $t > gpurun_out/ncu_tc_$t.log 2>&1; echo "ncu $t exit $?"
done
timeout 900 python benchmarks/bench_ops.py --sizes 1024,2048,4096,8192,16384,32768,65536,100000 --out gpurun_out/ops_y.jsonl > /dev/null 2> gpurun_out/ops_y.err; echo "ops exit $?"; tail -3 gpurun_out/ops_y.err
