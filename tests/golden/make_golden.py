"""Generates tests/golden/*.npz ON THE GPU BOX from the reference's own CUDA kernels
(oracle/_ref/libpointnet2_ref.so = /root/reference/openpoints/cpp/pointnet2_batch/src/*_gpu.cu compiled
unmodified).  Run:  gpurun -- python tests/golden/make_golden.py   then copy gpurun_out/golden/*.npz here."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from adaptpoint_b200.synth import make_clouds  # noqa: E402
from oracle import refcuda  # noqa: E402

CASES = [("quant", 2, 600, 150, 0.25, 16), ("dup", 2, 1024, 256, 0.2, 32), ("shell", 1, 1500, 300, 0.15, 24),
         ("same", 1, 40, 8, 0.1, 4), ("quant", 2, 96, 96, 0.5, 8)]


def main():
    dev = torch.device("cuda:0")
    out_dir = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for i, (kind, B, N, m, radius, K) in enumerate(CASES):
        xyz = make_clouds(B, N, kind, seed=100 + i)
        xd = xyz.to(dev)
        fps_idx = refcuda.fps(xd, m)
        query = torch.gather(xd, 1, fps_idx.long().unsqueeze(-1).expand(-1, -1, 3)).contiguous()
        bq_idx = refcuda.ball_query(radius, K, xd, query)
        nn_d2, nn_idx = refcuda.three_nn(xd, query)
        g = torch.Generator().manual_seed(i)
        feat = torch.randn(B, 6, N, generator=g)
        feat_known = torch.randn(B, 5, m, generator=g)
        w = torch.rand(B, N, 3, generator=g)
        group_out = refcuda.group(feat.to(dev), bq_idx)
        interp_out = refcuda.three_interp(feat_known.to(dev), nn_idx, w.to(dev))
        np.savez_compressed(
            os.path.join(out_dir, f"case{i}_{kind}_N{N}.npz"), xyz=xyz.numpy(), fps_m=m, fps_idx=fps_idx.cpu().numpy(),
            bq_radius=np.float32(radius), bq_k=K, bq_query=query.cpu().numpy(), bq_idx=bq_idx.cpu().numpy(),
            nn_d2=nn_d2.cpu().numpy(), nn_idx=nn_idx.cpu().numpy(), feat=feat.numpy(), feat_known=feat_known.numpy(),
            interp_w=w.numpy(), group_out=group_out.cpu().numpy(), interp_out=interp_out.cpu().numpy())
        print("wrote case", i, kind)


if __name__ == "__main__":
    main()
