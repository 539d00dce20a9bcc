set -x
mkdir -p gpurun_out
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --dump-kernels gpurun_out/kernels_r01_ad.json > gpurun_out/bench_r01_ad.json 2> gpurun_out/bench_r01_ad.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_ad.json; tail -3 gpurun_out/bench_r01_ad.err
