"""Which Python lines launch the step's torch (non-libapb200) kernels?  One eager AdaptPoint step under torch.profiler;
every top-level aten op that launched CUDA kernels is attributed to the innermost frame inside adaptpoint_b200/ (forward)
or to its autograd node (backward).  Run on the GPU box:  python scripts/profile_torch_ops.py > gpurun_out/torch_ops.txt"""
import collections
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import loader_batch, B_PER_GPU, N_POINTS  # noqa: E402
from adaptpoint_b200.train import AdaptPointTrainer  # noqa: E402

dev = torch.device("cuda:0")
tr = AdaptPointTrainer(device=dev, seed=0)
pts, y = loader_batch(B_PER_GPU, N_POINTS, 100)
pts, y = pts.to(dev), y.to(dev)
for _ in range(3):
    tr.step(pts.clone(), y)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
    tr.step(pts.clone(), y)
    torch.cuda.synchronize()

by_site = collections.defaultdict(lambda: [0, 0.0, collections.Counter()])
total_k, total_us = 0, 0.0
for ev in prof.events():
    if not ev.kernels:
        continue
    par = ev.cpu_parent
    if par is not None and par.kernels and par.name.startswith("aten::"):
        continue                      # counted with its aten parent
    nk = len(ev.kernels)
    us = sum(k.duration for k in ev.kernels)
    site = None
    for fr in (ev.stack or []):
        if "adaptpoint_b200/" in fr:
            site = fr.split("adaptpoint_b200/")[-1].strip()
            break
    if site is None:
        p = ev
        while p is not None:
            if "evaluate_function" in p.name or "Backward" in p.name:
                site = "autograd: " + p.name.replace("autograd::engine::evaluate_function: ", "")
                break
            p = p.cpu_parent
    if site is None:
        p, top = ev, ev
        while p is not None:
            top = p
            p = p.cpu_parent
        site = "other: " + top.name
    rec = by_site[site]
    rec[0] += nk
    rec[1] += us
    rec[2][ev.name] += nk
    total_k += nk
    total_us += us
print(f"kernels launched in one eager step: {total_k}, summed kernel time {total_us / 1e3:.2f} ms")
for site, (nk, us, ops) in sorted(by_site.items(), key=lambda kv: -kv[1][0])[:70]:
    top = ", ".join(f"{n.replace('aten::', '')}x{c}" for n, c in ops.most_common(4))
    print(f"{nk:5d} kernels {us / 1e3:7.3f} ms  {site[:90]:90s} {top}")
