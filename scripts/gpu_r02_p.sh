# Round 2, call p: ncu full captures of the kernels changed this round + launch list of the final step
set -x
mkdir -p gpurun_out
cap() {  # name regex target args...
  name=$1; rx=$2; shift 2
  timeout 120 python scripts/ncu_targets.py "$@" > gpurun_out/ncu_plain_$name.log 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$rx -s 2 -c 1 -f -o gpurun_out/ncu_r02_p_$name python scripts/ncu_targets.py "$@" > gpurun_out/ncu_r02_p_$name.log 2>&1; echo "ncu $name exit $?"
}
cap wgrad tc_wgrad wgrad 32 256 512 2048
cap attn_fwd attn_mma_fwd attn
cap attn_dkv attn_mma_bwd_dkv attn
cap ystats group_ystats la
cap dense sa1_bwd_dense la
cap pmax tc_cm_gemm pmax
timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_step.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/launches_r02_p.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_step.log 2>&1; echo "ncu step exit $?"
