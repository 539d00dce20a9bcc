set -x
mkdir -p gpurun_out
timeout 200 python bench.py --steps 10 --warmup 3 --dump-kernels gpurun_out/kernels_r01_al.json > gpurun_out/bench_r01_al.json 2> gpurun_out/bench_r01_al.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_r01_al.json; tail -3 gpurun_out/bench_r01_al.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/launches_r01_al.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_al.log 2>&1; echo "ncu list exit $?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_dkv -c 1 -f -o gpurun_out/attn_dkv_full_al python -c "
import torch, sys
sys.path.insert(0, '.')
from adaptpoint_b200 import ops
q = torch.randn(32, 2048, 64, device='cuda', requires_grad=True)
k = torch.randn(32, 2048, 64, device='cuda', requires_grad=True)
v = torch.randn(32, 2048, 64, device='cuda', requires_grad=True)
for _ in range(2):
    o = ops.self_attention(q, k, v, 4)
    o.sum().backward()
torch.cuda.synchronize()
" > gpurun_out/ncu_attn.log 2>&1; echo "ncu attn exit $?"
