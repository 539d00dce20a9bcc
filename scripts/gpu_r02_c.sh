set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_geometry_gpu.py -q 2>&1 | tail -12 | cut -c1-300
