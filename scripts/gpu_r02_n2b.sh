# 2-GPU call + a single-GPU re-measurement on the same box (N=1 and N=2 from one machine)
set -x
mkdir -p gpurun_out
bash scripts/gpu_r02_scale.sh 2
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_r02_final5_n1b.json 2> /dev/null; echo "bench N=1 exit $?"; cut -c1-260 gpurun_out/bench_r02_final5_n1b.json
CUDA_VISIBLE_DEVICES=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_r02_final5_n1c.json 2> /dev/null; echo "bench N=1 (gpu 1) exit $?"; cut -c1-260 gpurun_out/bench_r02_final5_n1c.json
