# Round 2, call a (one B200): the whole -m gpu suite (without -x: collect every failure of the new tc-profile cases),
# smoke(), the contract bench through the new teardown, graph-captured cfg4/cfg5 model steps left unmeasured in round 1.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -s > gpurun_out/r02_a_tests.log 2>&1; echo "tests exit $?"; grep -E "passed|failed" gpurun_out/r02_a_tests.log | tail -3
grep -E "^FAILED|^ERROR" gpurun_out/r02_a_tests.log | head -40
grep "step parity" gpurun_out/r02_a_tests.log | cut -c1-1500
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_a_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r02_a_smoke.log
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r02_a.json 2> gpurun_out/bench_r02_a.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_r02_a.json; tail -3 gpurun_out/bench_r02_a.err
for c in b-cls xl-seg; do
  timeout 150 python benchmarks/bench_models.py --steps 10 --warmup 3 --cases $c --graph >> gpurun_out/models_r02_a_graph_n1.jsonl 2> gpurun_out/models_r02_a_graph_$c.err; echo "$c graph exit $?"; tail -3 gpurun_out/models_r02_a_graph_$c.err
done
cut -c1-330 gpurun_out/models_r02_a_graph_n1.jsonl
