set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_attention_gpu.py -q > gpurun_out/r02_g_attn.log 2>&1; echo "attn tests exit $?"; tail -12 gpurun_out/r02_g_attn.log | cut -c1-250
timeout 120 python scripts/bench_attn.py > gpurun_out/attn_r02_g.json 2>&1; echo "attn bench exit $?"; cat gpurun_out/attn_r02_g.json
timeout 400 python bench.py --steps 20 --warmup 5 --no-cpu --dump-kernels gpurun_out/kernels_r02_g.json > gpurun_out/bench_r02_g.json 2> gpurun_out/bench_r02_g.err; echo "bench exit $?"; cut -c1-240 gpurun_out/bench_r02_g.json; tail -3 gpurun_out/bench_r02_g.err
