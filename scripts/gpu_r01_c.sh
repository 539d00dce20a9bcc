set -x
mkdir -p gpurun_out
python benchmarks/bench_sa.py --out gpurun_out/sa_r01_c.jsonl > gpurun_out/sa_r01_c.log 2>&1; echo "sa exit $?"; tail -5 gpurun_out/sa_r01_c.log
python scripts/ncu_targets.py sa2 && ncu --set full --clock-control none --import-source on -k regex:tc_rows_gemm -c 2 -o gpurun_out/ncu_sa2_r01 -f python scripts/ncu_targets.py sa2 > gpurun_out/ncu_sa2.log 2>&1; echo "ncu sa2 exit $?"
python scripts/ncu_targets.py pw && ncu --set full --clock-control none --import-source on -k regex:tc_rows_gemm -c 1 -o gpurun_out/ncu_pw_r01 -f python scripts/ncu_targets.py pw > gpurun_out/ncu_pw.log 2>&1; echo "ncu pw exit $?"
ls -la gpurun_out
