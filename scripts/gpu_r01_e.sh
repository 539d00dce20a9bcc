set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tcconv_gpu.py tests/test_train_gpu.py -x -q 2>&1 | tail -15
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r01_d.json 2> gpurun_out/bench_r01_d.err; echo "bench exit $?"; cut -c1-1200 gpurun_out/bench_r01_d.json; tail -5 gpurun_out/bench_r01_d.err
python bench.py --steps 10 --warmup 3 --no-graph --no-cpu > gpurun_out/bench_r01_d_nograph.json 2> gpurun_out/bench_r01_d_ng.err; echo "bench exit $?"; cut -c1-400 gpurun_out/bench_r01_d_nograph.json; tail -5 gpurun_out/bench_r01_d_ng.err
