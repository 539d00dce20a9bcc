"""Sweep the slice size of the sliced grouping kernels (rows longer than shared memory): python scripts/seg_sweep.py"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from adaptpoint_b200 import _cabi, ops  # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
from bench_ops import timeit  # noqa: E402

dev = torch.device("cuda:0")
flush = torch.zeros(160 * 1024 * 1024 // 4, device=dev)
B, K = 32, 32
for N in (65536, 100000):
    M = N // 4
    nbr = torch.randint(0, N, (B, M, K), device=dev, dtype=torch.int32)
    for C in (32,):
        feat = torch.randn(B, C, N, device=dev)
        go = torch.randn(B, C, M, K, device=dev)
        gp = torch.zeros(B, C, N, device=dev)

        def bwd():
            gp.zero_()
            _cabi.call("apb_group_grad_f32", [go, nbr, gp, B, C, N, M, K], dev)
        for kb in (200, 136, 100, 66, 50, 32):
            _cabi.tune_group_seg(kb)
            f, _ = timeit(lambda: ops.grouping_operation(feat, nbr), 5, flush=flush)
            b, _ = timeit(bwd, 5, flush=flush)
            print(json.dumps({"N": N, "C": C, "seg_kb": kb, "fwd_ms": round(f, 4), "bwd_ms": round(b, 4),
                              "fwd_frac": round(4.0 * B * (C * N + M * K + C * M * K) / 1e9 / (f / 1e3) / 6539.2, 3),
                              "bwd_frac": round(4.0 * B * (C * N + M * K + C * M * K) / 1e9 / (b / 1e3) / 6539.2, 3)}),
                  flush=True)
        _cabi.tune_group_seg(0)
