# Round 2, call d: whole GPU suite, contract bench, ncu launch list of the step, ncu full captures of the TMA-fed GEMM.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/r02_d_tests.log 2>&1; echo "tests exit $?"; grep -E "passed|failed" gpurun_out/r02_d_tests.log | tail -2; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02_d_tests.log | head -30
timeout 400 python bench.py --steps 20 --warmup 5 --dump-kernels gpurun_out/kernels_r02_d.json > gpurun_out/bench_r02_d.json 2> gpurun_out/bench_r02_d.err; echo "bench exit $?"; cut -c1-240 gpurun_out/bench_r02_d.json; tail -3 gpurun_out/bench_r02_d.err
i=0
for shape in "32 128 512 2048" "8 512 512 65536" "8 64 256 24000"; do
  i=$((i+1))
  timeout 120 python scripts/ncu_targets.py pwtf32 $shape > gpurun_out/ncu_plain_$i.log 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:tc_cm_gemm -s 2 -c 1 -f -o gpurun_out/ncu_r02_d_pw_$i python scripts/ncu_targets.py pwtf32 $shape > gpurun_out/ncu_r02_d_pw_$i.log 2>&1; echo "ncu pw $i exit $?"
done
timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_step.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/launches_r02_d.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_step.log 2>&1; echo "ncu step exit $?"
ls -la gpurun_out | tail -15
