"""Shared-MLP microbench: the tcgen05 kernels of csrc/sa_fused.cu vs the unfused fp32 op sequence (the
reference's: ball query -> grouping x2 -> cat -> Conv2d/BN/ReLU -> max) on PointNeXt-S stage shapes (B=32) and
PointNeXt-XL S3DIS-shaped LocalAggregation / pwconv shapes.  Reports time, executed tensor FLOP/s against the
measured bf16 peak, and the speed-up.  JSON lines to stdout / --out."""
import argparse
import json
import os
import sys

import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from adaptpoint_b200 import _cabi, fused, ops  # noqa: E402
from adaptpoint_b200.layers import create_convblock1d, get_aggregation_feautres  # noqa: E402
from adaptpoint_b200.pointnext import LocalAggregation  # noqa: E402
from adaptpoint_b200.synth import make_clouds  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["bf16_tflops"], d["hbm_gbs"], "measured"
    return 1800.0, 6650.0, "fallback"


def timeit(fn, iters=10, warm=3, flush=None):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.add_(1.0)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def kernel_times(fn):
    """Per-launch CUDA-event times of the C-ABI calls made by fn (one traced invocation)."""
    _cabi.TRACE = []
    fn()
    torch.cuda.synchronize()
    tr, _cabi.TRACE = _cabi.TRACE, None
    return {name: e0.elapsed_time(e1) for name, _, e0, e1 in tr}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--B", type=int, default=32)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    tf_peak, hbm, src = peaks()
    flush = torch.zeros(160 * 1024 * 1024 // 4, device=dev)
    lines = []

    def emit(d):
        lines.append(d)
        print(json.dumps(d), flush=True)

    B = args.B
    # (name, B, N, M, cin, channels, radius)
    sa_cases = [("S.stage1", B, 1024, 512, 32, [32, 64], 0.15), ("S.stage2", B, 512, 256, 64, [64, 128], 0.225),
                ("S.stage3", B, 256, 128, 128, [128, 256], 0.3375), ("S.stage4", B, 128, 64, 256, [256, 512], 0.50625),
                ("S.stage1.N2048", B, 2048, 1024, 32, [32, 64], 0.15),
                ("XL.sa1", 8, 24000, 6000, 64, [128], 0.1), ("XL.la1", 8, 6000, 6000, 128, [128], 0.2),
                ("XL.la2", 8, 1500, 1500, 256, [256], 0.4), ("XL.la3", 8, 375, 375, 512, [512], 0.8),
                ("wide2.N8192", 8, 8192, 2048, 128, [128, 256], 0.1)]
    for name, b, N, M, cin, ch, r in sa_cases:
        la = LocalAggregation([cin] + ch, norm_args={"norm": "bn"}, act_args={"act": "relu"},
                              group_args={"NAME": "ballquery", "radius": r, "nsample": 32, "normalize_dp": True},
                              last_act=len(ch) == 1).to(dev).eval()
        xyz = make_clouds(b, N, "shell", 0).to(dev)
        f = torch.randn(b, cin, N, device=dev)
        if M == N:
            q = xyz
        else:
            ix = ops.furthest_point_sample(xyz, M).long()
            q = torch.gather(xyz, 1, ix.unsqueeze(-1).expand(-1, -1, 3)).contiguous()

        def unfused():
            with torch.no_grad():
                dp, fj = la.grouper(q, xyz, f)
                return la.pool(la.convs(get_aggregation_feautres(q, dp, None, fj, "dp_fj")))

        def fusedfn():
            return fused.sa_fused_forward(la, la.grouper, la.convs, q, xyz, f)

        ms_u = timeit(unfused, 5, 2, flush)
        ms_f = timeit(fusedfn, 10, 3, flush)
        kt = kernel_times(fusedfn)
        K = 32
        k1p = (cin + 63) // 64 * 64
        flops_u = 2.0 * b * N * k1p * ch[0]
        d = {"case": name, "B": b, "N": N, "M": M, "K": K, "cin": cin, "channels": ch, "unfused_fp32_ms": round(ms_u, 4),
             "fused_ms": round(ms_f, 4), "speedup": round(ms_u / ms_f, 2),
             "ball_query_ms": round(kt.get("apb_ball_query_f32", 0.0), 4),
             "sa_fused_call_ms": round(kt.get("apb_sa_fused_fwd_bf16", 0.0), 4),
             "ref_flops_G": round(2.0 * b * M * K * ((cin + 3) * ch[0] + (ch[0] * ch[1] if len(ch) == 2 else 0)) / 1e9, 3)}
        if len(ch) == 2:
            k2p = (ch[0] + 63) // 64 * 64
            flops2 = 2.0 * b * M * K * k2p * ((ch[1] + 15) // 16 * 16)
            d["executed_tensor_GF"] = round((flops_u + flops2) / 1e9, 3)
            d["tensor_TFs_over_call"] = round((flops_u + flops2) / 1e12 / (d["sa_fused_call_ms"] / 1e3), 2)
            d["tensor_frac_of_bf16_peak"] = round(d["tensor_TFs_over_call"] / tf_peak, 4)
        else:
            d["executed_tensor_GF"] = round(flops_u / 1e9, 3)
        # minimal HBM bytes of the fused call: features + xyz + idx in, U out+in (L2), out
        cout = ch[-1]
        d["min_bytes_MB"] = round(4.0 * b * (cin * N + 3 * N + 3 * M + M * K + cout * M + 2 * N * ch[0]) / 1e6, 2)
        d["min_bytes_GBs"] = round(d["min_bytes_MB"] / 1e3 / (d["sa_fused_call_ms"] / 1e3), 1)
        d["hbm_frac"] = round(d["min_bytes_GBs"] / hbm, 4)
        emit(d)
        del la, f

    # plain 1x1 conv (+BN+ReLU): (name, B, cin, cout, P)
    pw_cases = [("S.conv2.stage1 grouped", B, 32, 64, 512 * 32), ("S.conv2.stage4 grouped", B, 256, 512, 64 * 32),
                ("disc 128->1024", B, 128, 1024 // 2, 2048), ("XL.pwconv 256->1024(512)", 8, 256, 512, 1500),
                ("XL.pwconv 512->512", 8, 512, 512, 6000), ("big 512->512 P=65536", 8, 512, 512, 65536),
                ("big 256->256 P=262144", 4, 256, 256, 262144)]
    for name, b, cin, cout, P in pw_cases:
        blk = create_convblock1d(cin, cout, norm_args={"norm": "bn1d"}, act_args={"act": "relu"}).to(dev).eval()
        x = torch.randn(b, cin, P, device=dev)

        def lib():
            with torch.no_grad():
                return blk(x)

        def ours():
            return fused.pwconv_forward(blk, x)

        ms_l = timeit(lib, 5, 2, flush)
        ms_o = timeit(ours, 10, 3, flush)
        kp = (cin + 63) // 64 * 64
        fl = 2.0 * b * P * kp * cout
        by = 4.0 * b * P * (cin + cout)
        emit({"case": "pwconv " + name, "B": b, "cin": cin, "cout": cout, "P": P, "torch_fp32_ms": round(ms_l, 4),
              "ours_bf16_ms": round(ms_o, 4), "speedup": round(ms_l / ms_o, 2),
              "tensor_TFs": round(fl / 1e12 / (ms_o / 1e3), 2), "tensor_frac_of_bf16_peak": round(fl / 1e12 / (ms_o / 1e3) / tf_peak, 4),
              "alg_GBs": round(by / 1e9 / (ms_o / 1e3), 1), "hbm_frac": round(by / 1e9 / (ms_o / 1e3) / hbm, 4),
              "peak_src": src})
        del blk, x
    if args.out:
        os.makedirs(os.path.dirname(args.out), exist_ok=True)
        with open(args.out, "w") as fo:
            for d in lines:
                fo.write(json.dumps(d) + "\n")


if __name__ == "__main__":
    main()
