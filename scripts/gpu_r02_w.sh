set -x
mkdir -p gpurun_out
timeout 300 python scripts/bench_sa2.py --out gpurun_out/sa2_microbench_r02_w.jsonl > gpurun_out/sa2_w.log 2>&1; echo "bench exit $?"; cut -c1-220 gpurun_out/sa2_w.log | tail -45
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sa2_rows -s 3 -c 1 -o gpurun_out/ncu_sa2_bwd_s0 python scripts/bench_sa2.py --only bwd --stage 0 --iters 1 > gpurun_out/ncu_w0.log 2>&1; echo "ncu exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sa2_rows -s 3 -c 1 -o gpurun_out/ncu_sa2_bwd_s3 python scripts/bench_sa2.py --only bwd --stage 3 --iters 1 > gpurun_out/ncu_w3.log 2>&1; echo "ncu exit $?"
