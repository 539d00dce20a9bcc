set -x
mkdir -p gpurun_out
timeout 300 python benchmarks/bench_models.py --steps 5 --warmup 3 --cases b-cls,xl-seg --dump-kernels gpurun_out/mk_r02_m > gpurun_out/models_r02_m.jsonl 2> gpurun_out/models_r02_m.err; echo "exit $?"; cut -c1-200 gpurun_out/models_r02_m.jsonl; tail -2 gpurun_out/models_r02_m.err
