set -x
mkdir -p gpurun_out
bash scripts/gpu_r02_scale.sh 2
timeout 300 python scripts/profile_torch_ops.py > gpurun_out/torch_ops_r02.txt 2> gpurun_out/torch_ops_r02.err; echo "profile exit $?"; head -3 gpurun_out/torch_ops_r02.txt
