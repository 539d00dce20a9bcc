"""One-conv set-abstraction blocks under autograd without the (B,C,M,K) tensors (fused_train.py, csrc/sa_train.cu:
U = Wf.f per support point, one gather pass for max/min over K + batch statistics, closed-form dense BatchNorm backward)
against the unfused op sequence of the reference (pointnext.py:66-77, :164-166) on the strict fp32 torch backend.
Two comparisons per case:
  exact : the three dense contractions of the fused path evaluated in fp32 (fused_train.EXACT_GEMM, test-only), so every
          kernel of csrc/sa_train.cu and the closed-form BatchNorm backward are checked ELEMENT BY ELEMENT at fp32
          tolerances (1e-4 outputs, 1e-3 gradients);
  tf32  : the product configuration (TMA-fed TF32 GEMMs): 3e-3 on outputs; gradients by direction and norm, because
          TF32 noise flips a fraction of the max-over-K routes and moves whole entries between support points."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-12)).item()


def _cos(a, b):
    return torch.nn.functional.cosine_similarity(a.double().flatten(), b.double().flatten(), dim=0).item()


def _inputs(dev, B, N, cin, seed, kind="ball"):
    from adaptpoint_b200.synth import make_clouds
    gen = torch.Generator().manual_seed(seed)
    p = make_clouds(B, N, kind, seed).to(dev)
    f = torch.randn(B, cin, N, generator=gen).to(dev)
    return p, f


def _run(block, p, f, backend, fused, gy_seed, stride=1, pos_grad=False, exact=False):
    from adaptpoint_b200 import fused_train, tcconv
    import _model_cases as mc
    mc.strict_fp32()
    fused_train.ENABLED = fused
    fused_train.EXACT_GEMM = exact
    pg = p.clone().requires_grad_(pos_grad)
    fg = f.clone().requires_grad_(True)
    try:
        with tcconv.conv_backend(backend):
            out = block([pg, fg])
            out = out[1] if isinstance(out, (list, tuple)) else out
            gy = torch.randn(out.shape, generator=torch.Generator().manual_seed(gy_seed)).to(out.device)
            out.backward(gy)
    finally:
        fused_train.ENABLED = True
        fused_train.EXACT_GEMM = False
    grads = {n: q.grad.detach().clone() for n, q in block.named_parameters() if q.grad is not None}
    bufs = {n: b.detach().clone() for n, b in block.named_buffers()}
    return out.detach(), fg.grad.detach(), (pg.grad.detach() if pos_grad else None), grads, bufs


def _compare_exact(got, want, pos_grad):
    assert _rel(got[0], want[0]) < 1e-4
    assert _rel(got[1], want[1]) < 1e-3, "df"
    if pos_grad:
        assert _rel(got[2], want[2]) < 1e-3, "dpos"
    for n, g in want[3].items():
        assert _rel(got[3][n], g) < 1e-3, n
    for n, b in want[4].items():
        if b.dtype.is_floating_point:
            assert _rel(got[4][n], b) < 1e-4, n
        else:
            assert torch.equal(got[4][n], b), n


def _compare_tf32(got, want, pos_grad):
    assert _rel(got[0], want[0]) < 3e-3
    assert _cos(got[1], want[1]) > 0.99 and abs(got[1].norm().item() / want[1].norm().item() - 1) < 2e-2, "df"
    if pos_grad:
        assert _cos(got[2], want[2]) > 0.99, "dpos"
    for n, g in want[3].items():
        assert _cos(got[3][n], g) > 0.995, n
        assert abs(got[3][n].norm().item() / g.norm().item() - 1) < 2e-2, n
    for n, b in want[4].items():
        if b.dtype.is_floating_point:
            assert _rel(got[4][n], b) < 3e-3, n


@pytest.mark.parametrize("B,N,cin,K,radius,training", [(4, 1024, 32, 32, 0.15, True), (2, 2048, 64, 32, 0.1, True),
                                                       (3, 512, 128, 16, 0.3, True), (2, 1024, 32, 32, 0.15, False)])
def test_local_aggregation_fused_train_vs_unfused(cuda_dev, B, N, cin, K, radius, training):
    from adaptpoint_b200.pointnext import LocalAggregation
    torch.manual_seed(cin + N)
    la = LocalAggregation([cin, cin], norm_args={"norm": "bn"}, act_args={"act": "relu"},
                          group_args={"NAME": "ballquery", "radius": radius, "nsample": K, "normalize_dp": True}).to(cuda_dev)
    with torch.no_grad():
        bn = la.convs[0][1]
        bn.weight.copy_(torch.randn(cin) * 0.7 + 0.3)            # negative gammas included (min instead of max route)
        bn.bias.copy_(torch.randn(cin) * 0.2)
        bn.running_mean.copy_(torch.randn(cin) * 0.1)
        bn.running_var.copy_(torch.rand(cin) + 0.5)
    la.train(training)
    p, f = _inputs(cuda_dev, B, N, cin, 7)
    ref, la2 = copy.deepcopy(la), copy.deepcopy(la)
    want = _run(ref, p, f, "torch", False, 3, pos_grad=not training)
    _compare_exact(_run(la2, p, f, "tc", True, 3, pos_grad=not training, exact=True), want, not training)
    _compare_tf32(_run(la, p, f, "tc", True, 3, pos_grad=not training), want, not training)


def test_set_abstraction_one_layer_fused_train_vs_unfused(cuda_dev):
    """Stride-4 SetAbstraction with sa_layers = 1 (PointNeXt-B/L/XL): FPS + query set != support set."""
    from adaptpoint_b200.pointnext import SetAbstraction
    torch.manual_seed(5)
    sa = SetAbstraction(32, 64, layers=1, stride=4, group_args={"NAME": "ballquery", "radius": 0.2, "nsample": 32,
                                                                "normalize_dp": True},
                        norm_args={"norm": "bn"}, act_args={"act": "relu"}).to(cuda_dev).train()
    p, f = _inputs(cuda_dev, 4, 2048, 32, 11, "shell")
    ref, sa2 = copy.deepcopy(sa), copy.deepcopy(sa)
    want = _run(ref, p, f, "torch", False, 9)
    _compare_exact(_run(sa2, p, f, "tc", True, 9, exact=True), want, False)
    _compare_tf32(_run(sa, p, f, "tc", True, 9), want, False)


def test_fused_train_never_builds_grouped_tensor(cuda_dev):
    """Peak memory of a LocalAggregation training pass stays far below one (B,C,N,K) activation."""
    from adaptpoint_b200 import tcconv
    from adaptpoint_b200.pointnext import LocalAggregation
    B, N, C, K = 8, 8192, 64, 32
    la = LocalAggregation([C, C], norm_args={"norm": "bn"}, act_args={"act": "relu"},
                          group_args={"NAME": "ballquery", "radius": 0.1, "nsample": K, "normalize_dp": True}).to(cuda_dev).train()
    p, f = _inputs(cuda_dev, B, N, C, 3)
    f.requires_grad_(True)
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats()
    base = torch.cuda.memory_allocated()
    with tcconv.conv_backend("tc"):
        out = la([p, f])
        out.sum().backward()
    torch.cuda.synchronize()
    peak = torch.cuda.max_memory_allocated() - base
    grouped = B * (C + 3) * N * K * 4
    assert peak < 0.5 * grouped, (peak, grouped)      # the unfused pass holds >= 5 such activations
