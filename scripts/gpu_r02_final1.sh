# Round 2 final single-GPU call: whole GPU suite, smoke, contract bench (both arms), cfg4/cfg5 model steps, cfg3 op sweep,
# sa2 microbenchmark, ncu full captures of the three sa_train2.cu kernels, launch list of the final step.
set -x
mkdir -p gpurun_out
T=r02_final
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/${T}_tests.log; grep -E "^FAILED|^E  " gpurun_out/${T}_tests.log | cut -c1-250 | head -10
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${T}_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 --dump-kernels gpurun_out/kernels_${T}_n1.json > gpurun_out/bench_${T}_n1.json 2> gpurun_out/bench_${T}_n1.err; echo "bench exit $?"; cut -c1-400 gpurun_out/bench_${T}_n1.json
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_${T}_ref_n1.json 2> /dev/null; echo "ref arm exit $?"; cut -c1-300 gpurun_out/bench_${T}_ref_n1.json
timeout 600 python benchmarks/bench_models.py --steps 10 --warmup 3 > gpurun_out/models_${T}_n1.jsonl 2> gpurun_out/models_${T}_n1.err; echo "models exit $?"; cut -c1-220 gpurun_out/models_${T}_n1.jsonl
timeout 300 python scripts/bench_sa2.py --out gpurun_out/sa2_microbench_${T}.jsonl > /dev/null 2>&1; echo "sa2 microbench exit $?"
timeout 900 python benchmarks/bench_ops.py --out gpurun_out/ops_microbench_${T}.jsonl > gpurun_out/ops_${T}.log 2>&1; echo "ops exit $?"
cap() {  # name regex skip args...
  name=$1; rx=$2; sk=$3; shift 3
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$rx -s $sk -c 1 -f -o gpurun_out/ncu_${T}_$name python scripts/bench_sa2.py "$@" > gpurun_out/ncu_${T}_$name.log 2>&1; echo "ncu $name exit $?"
}
cap sa2_fwd_s1 sa2_fwd 3 --only fwd --stage 1 --iters 1
cap sa2_bwd_s1 sa2_bwd 3 --only bwd --stage 1 --iters 1
cap sa2_wgrad_s1 sa2_wgrad 3 --only wgrad --stage 1 --iters 1
cap sa2_fwd_s3 sa2_fwd 3 --only fwd --stage 3 --iters 1
timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_step.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/launches_${T}.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_step.log 2>&1; echo "ncu step exit $?"
