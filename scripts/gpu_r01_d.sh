set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fused_gpu.py tests/test_tcconv_gpu.py -x -q 2>&1 | tail -25
python benchmarks/bench_sa.py --out gpurun_out/sa_r01_e.jsonl > gpurun_out/sa_r01_e.log 2>&1; echo "sa exit $?"
python scripts/ncu_targets.py sa2 && ncu --set full --clock-control none --import-source on -k regex:tc_rows_gemm -c 2 -o gpurun_out/ncu_sa2_r01b -f python scripts/ncu_targets.py sa2 > gpurun_out/ncu_sa2.log 2>&1; echo "ncu sa2 exit $?"
python scripts/ncu_targets.py pw && ncu --set full --clock-control none --import-source on -k regex:tc_rows_gemm -c 1 -o gpurun_out/ncu_pw_r01b -f python scripts/ncu_targets.py pw > gpurun_out/ncu_pw.log 2>&1; echo "ncu pw exit $?"
