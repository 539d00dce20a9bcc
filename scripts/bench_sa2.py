"""Microbenchmark of the two-conv fused set-abstraction kernels (csrc/sa_train2.cu) on the four PointNeXt-S stages at
B = 32 (the shapes of BASELINE configs[1]): forward, backward rows pass (training and eval mode), weight gradient.
   python scripts/bench_sa2.py [--out file.jsonl] [--only fwd|bwd|wgrad] [--stage i] [--iters n]
CUDA-event timing after warm-up, L2 flushed between iterations; one JSON line per (kernel, shape)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adaptpoint_b200 import _cabi, ops, tcconv      # noqa: E402
from adaptpoint_b200.synth import make_clouds       # noqa: E402

STAGES = [(2048, 1024, 32, 64, 0.15), (1024, 512, 64, 128, 0.3), (512, 256, 128, 256, 0.6), (256, 128, 256, 512, 1.2),
          (1024, 512, 32, 64, 0.15), (512, 256, 64, 128, 0.3), (256, 128, 128, 256, 0.6), (128, 64, 256, 512, 1.2)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--only", default="")
    ap.add_argument("--stage", type=int, default=-1)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--dbg", type=int, default=0, help="apb_tune_sa2 flags (timing experiments, wrong results)")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    _cabi.load().apb_tune_sa2(args.dbg)
    B, K = args.batch, 32
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    lines = []

    def timeit(fn):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(args.iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts.sort()
        return ts[len(ts) // 2]

    for si, (N, M, C1, C2, radius) in enumerate(STAGES):
        if args.stage >= 0 and si != args.stage:
            continue
        g = torch.Generator().manual_seed(si)
        p = make_clouds(B, N, "ball", si).to(dev)
        q = p[:, :M].contiguous()
        idx = ops.ball_query(radius, K, p, q).int().contiguous()
        u = torch.randn(B, N, C1, generator=g).to(dev)
        wp = torch.randn(C1, 3, generator=g).to(dev)
        s1 = (torch.randn(C1, generator=g) * 0.5 + 0.5).to(dev)
        t1 = (torch.randn(C1, generator=g) * 0.3).to(dev)
        w2 = (torch.randn(C2, C1, generator=g) / C1 ** 0.5).to(dev)
        w2b, k2p = tcconv.pack_f32(w2)
        vmax = torch.empty((B, C2, M), device=dev)
        vmin = torch.empty_like(vmax)
        arg = torch.empty((B, C2, M), dtype=torch.int32, device=dev)
        stats = torch.zeros((C2, 2), dtype=torch.float64, device=dev)
        gs_pm = torch.randn(B, M, C2, generator=g).to(dev)
        ksel_pm = torch.randint(0, K, (B, M, C2), generator=g, dtype=torch.uint8).to(dev)
        kc2, c1pad = (C2 + 31) // 32, (C1 + 31) // 32 * 32
        mean1, rstd1 = torch.zeros(C1, device=dev), torch.ones(C1, device=dev)
        du = torch.zeros((B, N, C1), device=dev)
        sums1 = torch.zeros((C1, 2), dtype=torch.float64, device=dev)
        sdzdp = torch.zeros((C1, 3), dtype=torch.float64, device=dev)
        sh1 = torch.zeros(C1, dtype=torch.float64, device=dev)
        dxyz, dq = torch.zeros_like(p), torch.zeros_like(q)
        rows = B * M * K
        base = {"stage": si, "B": B, "N": N, "M": M, "K": K, "C1": C1, "C2": C2, "rows": rows}

        def emit(kernel, us, flops, extra=None):
            d = dict(base, kernel=kernel, us=round(us, 1), tflops=round(flops / us / 1e6, 1))
            d.update(extra or {})
            lines.append(d)
            print(json.dumps(d), flush=True)

        if args.only in ("", "fwd"):
            us = timeit(lambda: _cabi.call("apb_sa2_fwd_tf32", [u, p, q, idx, wp, s1, t1, 1, w2b, vmax, vmin, arg, stats, B, N, M,
                                                               K, C1, C2, k2p], dev))
            emit("sa2_fwd", us, 2.0 * rows * C1 * C2)
        if args.only in ("", "bwd"):
            for training in (True, False):
                wb = torch.zeros((C1, kc2 * 32 + (c1pad if training else 0)), device=dev)
                wb[:, :C2] = w2.t()
                vvec = torch.zeros(C1, device=dev) if training else None
                a = [u, p, q, idx, wp, s1, t1, 1, gs_pm, ksel_pm, wb, wb.shape[1], kc2, vvec, mean1, rstd1, du, sums1, sdzdp, sh1,
                     None if training else dxyz, None if training else dq, B, N, M, K, C1, C2]
                us = timeit(lambda: _cabi.call("apb_sa2_bwd_rows_tf32", a, dev))
                emit("sa2_bwd_rows", us, 2.0 * rows * C1 * wb.shape[1], {"training": training})
        if args.only in ("", "wgrad"):
            for gram in (1, 0):
                dmat = torch.zeros((C2 + (C1 if gram else 0), c1pad), device=dev)
                us = timeit(lambda: _cabi.call("apb_sa2_wgrad_tf32", [u, p, q, idx, wp, s1, t1, 1, gs_pm, ksel_pm, dmat, gram, B, N,
                                                                     M, K, C1, C2], dev))
                emit("sa2_wgrad", us, 2.0 * rows * C1 * (C2 + (C1 if gram else 0)), {"gram": gram})
    if args.out:
        with open(args.out, "w") as f:
            for d in lines:
                f.write(json.dumps(d) + "\n")


if __name__ == "__main__":
    main()
