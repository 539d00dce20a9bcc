"""Forward GPU time per sub-module (CUDA events in forward hooks, eager, serial launch order) of the three networks
in one AdaptPoint step:  python scripts/profile_modules.py > gpurun_out/module_profile.txt"""
import os
import sys
from collections import defaultdict

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import loader_batch, B_PER_GPU, N_POINTS  # noqa: E402
from adaptpoint_b200.train import AdaptPointTrainer  # noqa: E402

dev = torch.device("cuda:0")
tr = AdaptPointTrainer(device=dev, seed=0, overlap=False, fused_eval=True)
pts, y = loader_batch(B_PER_GPU, N_POINTS, 100)
pts, y = pts.to(dev), y.to(dev)
for _ in range(3):
    tr.step(pts.clone(), y)
torch.cuda.synchronize()

records = []


def hook(net, depth):
    for name, m in net.named_modules():
        if name == "" or name.count(".") >= depth:
            continue
        tag = f"{type(net).__name__}.{name} [{type(m).__name__}]"

        def pre(mod, inp, tag=tag):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            mod._apb_e0 = e

        def post(mod, inp, out, tag=tag):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            records.append((tag, mod._apb_e0, e))
        m.register_forward_pre_hook(pre)
        m.register_forward_hook(post)


hook(tr.generator, 3)
hook(tr.discriminator, 2)
hook(tr.classifier, 3)
steps = 3
for _ in range(steps):
    tr.step(pts.clone(), y)
torch.cuda.synchronize()
tot, cnt = defaultdict(float), defaultdict(int)
for tag, e0, e1 in records:
    tot[tag] += e0.elapsed_time(e1)
    cnt[tag] += 1
print(f"forward GPU ms per step by module ({steps} steps, serial order)")
for tag, ms in sorted(tot.items(), key=lambda kv: -kv[1])[:70]:
    print(f"{ms / steps:8.3f} ms  {cnt[tag] // steps:3d} calls  {tag}")
