set -x
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_parity_gpu.py -x -q 2>&1 | tail -4
timeout 150 python scripts/fps_sweep.py --cases bench > gpurun_out/fps_sweep_bench.md 2> gpurun_out/fps_sweep.err; echo "exit $?"; tail -3 gpurun_out/fps_sweep.err; cat gpurun_out/fps_sweep_bench.md
timeout 300 python scripts/fps_sweep.py --cases sweep > gpurun_out/fps_sweep_big.md 2>> gpurun_out/fps_sweep.err; echo "exit $?"; tail -3 gpurun_out/fps_sweep.err; cat gpurun_out/fps_sweep_big.md
