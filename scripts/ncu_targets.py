"""Small single-op programs for `ncu --set full` captures:  python scripts/ncu_targets.py <target>
targets: sa2 (two-layer fused SA, PointNeXt-S stage 1, B=32), pw (1x1 conv 512->512), pwtf32, wgrad, attn, la (one-conv
LocalAggregation training step, fused_train), fps, group, group_grad"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from adaptpoint_b200 import _cabi, fused, ops  # noqa: E402
from adaptpoint_b200.layers import create_convblock1d  # noqa: E402
from adaptpoint_b200.pointnext import LocalAggregation  # noqa: E402
from adaptpoint_b200.synth import make_clouds  # noqa: E402

dev = torch.device("cuda:0")
tgt = sys.argv[1] if len(sys.argv) > 1 else "sa2"
if tgt == "sa2":
    B, N, M, cin, ch, r = 32, 2048, 1024, 32, [32, 64], 0.15
    la = LocalAggregation([cin] + ch, norm_args={"norm": "bn"}, act_args={"act": "relu"},
                          group_args={"NAME": "ballquery", "radius": r, "nsample": 32, "normalize_dp": True},
                          last_act=False).to(dev).eval()
    xyz = make_clouds(B, N, "shell", 0).to(dev)
    f = torch.randn(B, cin, N, device=dev)
    ix = ops.furthest_point_sample(xyz, M).long()
    q = torch.gather(xyz, 1, ix.unsqueeze(-1).expand(-1, -1, 3)).contiguous()
    for _ in range(3):
        fused.sa_fused_forward(la, la.grouper, la.convs, q, xyz, f)
elif tgt == "pw":
    blk = create_convblock1d(512, 512, norm_args={"norm": "bn1d"}, act_args={"act": "relu"}).to(dev).eval()
    x = torch.randn(8, 512, 65536, device=dev)
    for _ in range(3):
        fused.pwconv_forward(blk, x)
elif tgt == "pwtf32":   # the bench step's most expensive 1x1-conv shape: python scripts/ncu_targets.py pwtf32 [B Cin Cout P]
    from adaptpoint_b200 import tcconv
    B, cin, cout, P = (int(v) for v in sys.argv[2:6]) if len(sys.argv) > 5 else (32, 128, 512, 2048)
    x = torch.randn(B, cin, P, device=dev)
    w = torch.randn(cout, cin, device=dev) * 0.05
    for _ in range(3):
        tcconv._gemm_rows(x, w, None, B, cin, cout, P)
elif tgt == "pmax":   # the discriminator's conv -> max over the points head (B=32, 128 -> 1024, N=2048)
    from adaptpoint_b200 import tcconv
    conv = tcconv.TcConv2d(128, 1024, 1).to(dev)
    x = torch.randn(32, 128, 1, 2048, device=dev)
    with torch.no_grad():
        for _ in range(3):
            tcconv.conv_max(conv, x)
elif tgt == "wgrad":
    from adaptpoint_b200 import tcconv
    B, cin, cout, P = (int(v) for v in sys.argv[2:6]) if len(sys.argv) > 5 else (32, 256, 512, 2048)
    x = torch.randn(B, cin, P, device=dev)
    g = torch.randn(B, cout, P, device=dev)
    for _ in range(3):
        tcconv.wgrad(g, x, B, cin, cout, P)
elif tgt == "attn":
    q, k, v = [torch.randn(32, 2048, 64, device=dev, requires_grad=True) for _ in range(3)]
    for _ in range(2):
        o = ops.self_attention(q, k, v, 4)
        o.backward(torch.ones_like(o))
elif tgt == "la":     # XL stage-1 LocalAggregation (B=8, N=6000, C=128): forward + backward of the fused training path
    from adaptpoint_b200 import tcconv
    B, N, C = 8, 6000, 128
    la = LocalAggregation([C, C], norm_args={"norm": "bn"}, act_args={"act": "relu"},
                          group_args={"NAME": "ballquery", "radius": 0.2, "nsample": 32, "normalize_dp": True}).to(dev).train()
    xyz = make_clouds(B, N, "ball", 0).to(dev)
    f = torch.randn(B, C, N, device=dev, requires_grad=True)
    for _ in range(3):
        out = la([xyz, f])
        out.sum().backward()
elif tgt in ("fps", "group", "group_grad"):
    B, N, M, K, C = 32, 16384, 4096, 32, 32
    if len(sys.argv) > 3:  # python scripts/ncu_targets.py fps 2048 1024 [C]
        N, M = int(sys.argv[2]), int(sys.argv[3])
    if len(sys.argv) > 4:
        C = int(sys.argv[4])
    xyz = make_clouds(B, N, "shell", 0).to(dev)
    for _ in range(2):
        idx = ops.furthest_point_sample(xyz, M)
    if tgt != "fps":
        q = torch.gather(xyz, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3)).contiguous()
        nbr = ops.ball_query(0.1, K, xyz, q)
        feat = torch.randn(B, C, N, device=dev, requires_grad=True)
        for _ in range(2):
            g = ops.grouping_operation(feat, nbr)
            if tgt == "group_grad":
                g.backward(torch.ones_like(g))
torch.cuda.synchronize()
print("done", tgt)
