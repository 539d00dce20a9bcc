"""FPS work-split sweep on B200: for each (B,N,m) time every forced (threads, cluster) plan (APB_FPS_FORCE) and
the automatic one, check all plans return identical indices (and equal the CPU oracle on cloud 0 for small N).
usage: python scripts/fps_sweep.py [--cases bench|sweep] > profiles/fps_config_sweep_rNN.md"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from adaptpoint_b200 import ops  # noqa: E402
from adaptpoint_b200.synth import make_clouds  # noqa: E402
from oracle import cpu as orc  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cases", default="bench")
args = ap.parse_args()
dev = torch.device("cuda:0")
if args.cases == "bench":
    cases = [(32, 256, 128), (32, 512, 256), (32, 1024, 512), (32, 2048, 1024), (32, 2048, 1200), (32, 2048, 4)]
else:
    cases = [(32, n, n // 4) for n in (4096, 8192, 16384, 32768, 65536, 100000)] + [(8, 24000, 6000), (1, 100000, 25000)]


def timeit(fn, iters=5):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    return sorted(ts)[len(ts) // 2]


print("| B | N | m | plan (threads,cluster) | ms | us/round | equal |\n|---|---|---|---|---|---|---|")
for B, N, m in cases:
    xyz = make_clouds(B, N, "shell", seed=1).to(dev)
    os.environ.pop("APB_FPS_FORCE", None)
    ref = ops.furthest_point_sample(xyz, m)
    if N <= 16384:
        want = orc.fps(xyz[:1].cpu().numpy(), m)
        assert np.array_equal(ref[:1].cpu().numpy(), want), f"auto plan differs from oracle at {B},{N},{m}"
    t = timeit(lambda: ops.furthest_point_sample(xyz, m), 3 if N > 30000 else 7)
    print(f"| {B} | {N} | {m} | auto | {t:.4f} | {1e3 * t / max(1, m - 1):.3f} | oracle-ok |", flush=True)
    for nt in (128, 256, 512, 1024):
        for c in (1, 2, 4, 8):
            os.environ["APB_FPS_FORCE"] = f"{nt},{c}"
            try:
                got = ops.furthest_point_sample(xyz, m)
            except Exception:
                continue
            eq = bool(torch.equal(got, ref))
            t = timeit(lambda: ops.furthest_point_sample(xyz, m), 3)
            print(f"| {B} | {N} | {m} | {nt},{c} | {t:.4f} | {1e3 * t / max(1, m - 1):.3f} | {eq} |", flush=True)
    os.environ.pop("APB_FPS_FORCE", None)
