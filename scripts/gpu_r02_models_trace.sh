set -x
mkdir -p gpurun_out
timeout 600 python benchmarks/bench_models.py --steps 10 --warmup 3 --cases b-cls,xl-seg --dump-kernels gpurun_out/models_trace_r02_final > gpurun_out/models_trace.jsonl 2> gpurun_out/models_trace.err; echo "exit $?"; cut -c1-160 gpurun_out/models_trace.jsonl
