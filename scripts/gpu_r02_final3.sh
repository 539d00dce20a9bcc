set -x
mkdir -p gpurun_out
T=r02_final3
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --dump-kernels gpurun_out/kernels_${T}_n1.json > gpurun_out/bench_${T}_n1.json 2> gpurun_out/bench_${T}_n1.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_${T}_n1.json; tail -3 gpurun_out/bench_${T}_n1.err
