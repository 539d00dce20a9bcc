set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_g.json 2> gpurun_out/bench_r01_g.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_r01_g.json; tail -3 gpurun_out/bench_r01_g.err
python bench.py --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches_r01_g.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_g.log 2>&1; echo "ncu list exit $?"
python scripts/ncu_targets.py fps 2048 1024 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:fps_kernel -c 1 -f -o gpurun_out/fps_2048_full python scripts/ncu_targets.py fps 2048 1024 > gpurun_out/ncu_fps.log 2>&1; echo "ncu fps exit $?"
python scripts/ncu_targets.py group 16384 4096 128 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:group_rows -c 1 -f -o gpurun_out/group_16k_c128_full python scripts/ncu_targets.py group 16384 4096 128 > gpurun_out/ncu_group.log 2>&1; echo "ncu group exit $?"
python benchmarks/bench_ops.py --sizes 16384,65536,100000 --out gpurun_out/ops_g.jsonl > /dev/null 2> gpurun_out/ops_g.err; echo "ops exit $?"; cut -c1-200 gpurun_out/ops_g.jsonl
