set -x
mkdir -p gpurun_out
timeout 300 python scripts/seg_sweep.py > gpurun_out/seg_sweep_s.jsonl 2> gpurun_out/seg_sweep_s.err; echo "exit $?"; cat gpurun_out/seg_sweep_s.jsonl; tail -3 gpurun_out/seg_sweep_s.err
