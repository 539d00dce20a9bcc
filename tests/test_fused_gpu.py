"""Parity of the tcgen05 shared-MLP kernels (csrc/sa_fused.cu) against the unfused fp32 module path
(the reference's op sequence: group.py:235-255 + pointnext.py:164-166), tolerance 2e-2 relative to the output
scale (bf16 operands, fp32 accumulation -- BASELINE.json north_star)."""
import pytest
import torch
import torch.nn as nn

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("torch_backend_default")]

TOL = 2e-2


def _rel_err(got, want):
    return ((got - want).abs().max() / want.abs().max().clamp_min(1e-6)).item()


def _randomise_bn(mod, gen):
    for m in mod.modules():
        if isinstance(m, (nn.BatchNorm1d, nn.BatchNorm2d)):
            m.running_mean.copy_(torch.randn(m.num_features, generator=gen) * 0.2)
            m.running_var.copy_(torch.rand(m.num_features, generator=gen) + 0.5)
            m.weight.data.copy_(torch.randn(m.num_features, generator=gen))      # negative gammas included
            m.bias.data.copy_(torch.randn(m.num_features, generator=gen) * 0.1)


@pytest.mark.parametrize("B,cin,cout,P,relu", [(2, 4, 32, 1024, True), (3, 35, 64, 777, False),
                                               (2, 131, 128, 300, True), (1, 259, 256, 129, True),
                                               (2, 64, 512, 260, False), (1, 515, 40, 64, True),
                                               (8, 259, 256, 8192, True), (4, 512, 512, 20000, False)])
def test_pwconv_matches_fp32(cuda_dev, B, cin, cout, P, relu):
    from adaptpoint_b200 import fused
    from adaptpoint_b200.layers import create_convblock1d
    gen = torch.Generator().manual_seed(cin * 7 + cout)
    blk = create_convblock1d(cin, cout, norm_args={"norm": "bn1d"}, act_args={"act": "relu"} if relu else None)
    _randomise_bn(blk, gen)
    blk = blk.to(cuda_dev).eval()
    x = torch.randn(B, cin, P, generator=gen).to(cuda_dev)
    with torch.no_grad():
        want = blk(x)
    got = fused.pwconv_forward(blk, x)
    assert got.shape == want.shape
    assert _rel_err(got, want) < TOL


def test_pwconv_4d_and_bias(cuda_dev):
    from adaptpoint_b200 import fused
    from adaptpoint_b200.layers import create_convblock2d
    gen = torch.Generator().manual_seed(5)
    blk = create_convblock2d(19, 48, norm_args=None, act_args=None).to(cuda_dev).eval()   # conv bias, no norm
    x = torch.randn(2, 19, 50, 32, generator=gen).to(cuda_dev)
    with torch.no_grad():
        want = blk(x)
    got = fused.pwconv_forward(blk, x)
    assert _rel_err(got, want) < TOL


def _sa_case(cuda_dev, B, N, M, cin, channels, radius, seed, normalize_dp=True):
    from adaptpoint_b200 import ops
    from adaptpoint_b200.pointnext import LocalAggregation
    from adaptpoint_b200.synth import make_clouds
    gen = torch.Generator().manual_seed(seed)
    ga = {"NAME": "ballquery", "radius": radius, "nsample": 32, "normalize_dp": normalize_dp}
    la = LocalAggregation([cin] + channels, norm_args={"norm": "bn"}, act_args={"act": "relu"}, group_args=ga,
                          last_act=len(channels) == 1)
    _randomise_bn(la, gen)
    la = la.to(cuda_dev).eval()
    xyz = make_clouds(B, N, "shell", seed).to(cuda_dev)
    f = torch.randn(B, cin, N, generator=gen).to(cuda_dev)
    if M == N:
        q = xyz
    else:
        idx = ops.furthest_point_sample(xyz, M).long()
        q = torch.gather(xyz, 1, idx.unsqueeze(-1).expand(-1, -1, 3)).contiguous()
    return la, q, xyz, f


@pytest.mark.parametrize("B,N,M,cin,channels,radius", [
    (2, 1024, 512, 32, [32, 64], 0.15),      # PointNeXt-S stage 1
    (2, 512, 256, 64, [64, 128], 0.225),     # stage 2
    (3, 256, 128, 128, [128, 256], 0.3375),  # stage 3
    (2, 128, 64, 256, [256, 512], 0.50625),  # stage 4 (Cout 512: single-buffered accumulator)
    (32, 128, 64, 256, [256, 512], 0.50625), # same, several tiles per CTA with a 1-deep ring per producer group
    (32, 1024, 512, 32, [32, 64], 0.15),     # full-batch stage 1: 28 tiles per CTA
    (1, 1000, 250, 16, [24, 40], 0.2),       # ragged channel counts
    (1, 333, 333, 8, [16, 16], 0.3),         # B*M*K not a multiple of 128
])
def test_sa_fused_two_layer(cuda_dev, B, N, M, cin, channels, radius):
    from adaptpoint_b200 import fused
    from adaptpoint_b200.layers import get_aggregation_feautres
    la, q, xyz, f = _sa_case(cuda_dev, B, N, M, cin, channels, radius, seed=N + cin)
    with torch.no_grad():
        dp, fj = la.grouper(q, xyz, f)
        want = la.pool(la.convs(get_aggregation_feautres(q, dp, None, fj, "dp_fj")))
    assert fused.fusable_reason(la.grouper, la.convs, 32) is None
    got = fused.sa_fused_forward(la, la.grouper, la.convs, q, xyz, f)
    assert got.shape == want.shape
    assert _rel_err(got, want) < TOL


@pytest.mark.parametrize("B,N,M,cin,cout,radius,ndp", [(2, 2048, 512, 32, 64, 0.1, True),
                                                      (2, 600, 600, 64, 64, 0.2, True),
                                                      (1, 1500, 375, 128, 256, 0.4, False),
                                                      (1, 300, 75, 256, 512, 0.8, True)])
def test_sa_fused_one_layer(cuda_dev, B, N, M, cin, cout, radius, ndp):
    from adaptpoint_b200 import fused
    from adaptpoint_b200.layers import get_aggregation_feautres
    la, q, xyz, f = _sa_case(cuda_dev, B, N, M, cin, [cout], radius, seed=N + cout, normalize_dp=ndp)
    with torch.no_grad():
        dp, fj = la.grouper(q, xyz, f)
        want = la.pool(la.convs(get_aggregation_feautres(q, dp, None, fj, "dp_fj")))
    got = fused.sa_fused_forward(la, la.grouper, la.convs, q, xyz, f)
    assert _rel_err(got, want) < TOL


def test_fused_model_logits(cuda_dev):
    """End-to-end: PointNeXt-S eval logits, fused=True vs the unfused fp32 path, same state_dict."""
    from adaptpoint_b200.models import build_model
    from adaptpoint_b200.synth import make_clouds
    torch.manual_seed(0)
    ref = build_model("pointnext-s").to(cuda_dev)
    _randomise_bn(ref.cpu(), torch.Generator().manual_seed(1))
    ref = ref.to(cuda_dev).eval()
    fus = build_model("pointnext-s", fused=True).to(cuda_dev).eval()
    fus.load_state_dict(ref.state_dict())
    xyz = make_clouds(4, 1024, "shell", 3).to(cuda_dev)
    x = torch.cat([xyz, xyz[:, :, 1:2] - xyz[:, :, 1:2].amin(1, keepdim=True)], -1).transpose(1, 2).contiguous()
    with torch.no_grad():
        want = ref({"pos": xyz, "x": x})
        got = fus({"pos": xyz, "x": x})
    assert _rel_err(got, want) < 5e-2     # six stacked bf16 layers
