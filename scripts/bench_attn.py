"""Attention microbenchmark (B=32, M=2048, H=4, d=16: the imitator's mask branch): forward and backward time, clocks."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from adaptpoint_b200 import ops
from bench import ClockSampler
dev = torch.device("cuda:0")
B, M, H = 32, 2048, 4
q, k, v = [torch.randn(B, M, H * 16, device=dev, requires_grad=True) for _ in range(3)]
go = torch.randn(B, M, H * 16, device=dev)
sampler = ClockSampler(0); sampler.start()
def t(fn, n=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
fwd = t(lambda: ops.self_attention(q, k, v, H))
o = ops.self_attention(q, k, v, H)
bwd = t(lambda: torch.autograd.grad(o, (q, k, v), go, retain_graph=True))
flops_f = 4.0 * B * H * M * M * 16
print(json.dumps({"op": "self_attention", "B": B, "M": M, "H": H, "fwd_ms": round(fwd, 4), "bwd_ms": round(bwd, 4),
                  "fwd_TFLOPs": round(flops_f / fwd / 1e9, 1), "bwd_TFLOPs": round(2.5 * flops_f / bwd / 1e9, 1),
                  "clocks": sampler.stop()}))
