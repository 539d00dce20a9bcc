set -x
mkdir -p gpurun_out
timeout 120 python scripts/ncu_targets.py attn > gpurun_out/ncu_plain_attn.log 2>&1; echo "plain exit $?"
for k in attn_mma_fwd attn_mma_bwd_dq attn_mma_bwd_dkv; do
timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -f -o gpurun_out/ncu_r02_final_$k python scripts/ncu_targets.py attn > gpurun_out/ncu_r02_final_$k.log 2>&1; echo "ncu $k exit $?"
done
ls -la gpurun_out/*.ncu-rep
