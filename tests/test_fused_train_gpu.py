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


# ---- two-conv blocks (csrc/sa_train2.cu): every kernel against its torch restatement, then the module against the
# ---- unfused op sequence -------------------------------------------------------------------------------------------
def _sa2_inputs(dev, B, N, M, C1, C2, seed):
    import adaptpoint_b200.ops as ops
    from adaptpoint_b200.synth import make_clouds
    g = torch.Generator().manual_seed(seed)
    p = make_clouds(B, N, "ball", seed).to(dev)
    q = p[:, :M].contiguous()
    idx = ops.ball_query(0.3, 32, p, q).int().contiguous()
    u = torch.randn(B, N, C1, generator=g).to(dev)
    wp = (torch.randn(C1, 3, generator=g) * 2).to(dev)
    s1 = (torch.randn(C1, generator=g) * 0.5 + 0.5).to(dev)
    t1 = (torch.randn(C1, generator=g) * 0.3).to(dev)
    w2 = (torch.randn(C2, C1, generator=g) / C1 ** 0.5).to(dev)
    return g, p, q, idx, u, wp, s1, t1, w2


# the last three give every CTA several row tiles (pipeline stages and TMEM accumulators are reused many times)
SA2_SHAPES = [(3, 256, 5, 32, 64), (4, 512, 128, 64, 128), (2, 256, 64, 256, 512), (2, 128, 33, 20, 36),
              (2, 256, 64, 128, 256), (8, 1024, 512, 128, 256), (16, 1024, 512, 32, 64), (8, 512, 256, 256, 512)]


@pytest.mark.parametrize("B,N,M,C1,C2", SA2_SHAPES)
def test_sa2_fwd_kernel_vs_restatement(cuda_dev, B, N, M, C1, C2):
    import _sa2_emul as emul
    from adaptpoint_b200 import _cabi, tcconv
    _, p, q, idx, u, wp, s1, t1, w2 = _sa2_inputs(cuda_dev, B, N, M, C1, C2, C1 + M)
    K = 32
    w2b, k2p = tcconv.pack_f32(w2)
    outs = []
    for fn in (lambda *a: _cabi.call("apb_sa2_fwd_tf32", list(a), cuda_dev), emul.apb_sa2_fwd_tf32):
        vmax = torch.empty((B, C2, M), device=cuda_dev)
        vmin = torch.empty_like(vmax)
        arg = torch.empty((B, C2, M), dtype=torch.int32, device=cuda_dev)
        stats = torch.zeros((C2, 2), dtype=torch.float64, device=cuda_dev)
        fn(u, p, q, idx, wp, s1, t1, 1, w2b, vmax, vmin, arg, stats, B, N, M, K, C1, C2, k2p)
        outs.append((vmax, vmin, arg, stats))
    torch.cuda.synchronize()
    got, want = outs
    scale = want[0].abs().max().item()
    assert (got[0] - want[0]).abs().max().item() < 4e-3 * scale, "vmax"
    assert (got[1] - want[1]).abs().max().item() < 4e-3 * scale, "vmin"
    assert (got[2] == want[2]).float().mean().item() > 0.97, "arg"       # TF32 noise flips near-ties only
    n = B * M * K
    assert (got[3][:, 0] - want[3][:, 0]).abs().max().item() < 2e-3 * scale * n ** 0.5 + 1e-3 * want[3][:, 0].abs().max().item()
    assert _rel(got[3][:, 1], want[3][:, 1]) < 2e-3, "sum of squares"


@pytest.mark.parametrize("training", [True, False])
@pytest.mark.parametrize("B,N,M,C1,C2", SA2_SHAPES)
def test_sa2_bwd_rows_kernel_vs_restatement(cuda_dev, B, N, M, C1, C2, training):
    import _sa2_emul as emul
    from adaptpoint_b200 import _cabi
    g, p, q, idx, u, wp, s1, t1, w2 = _sa2_inputs(cuda_dev, B, N, M, C1, C2, C2 + M)
    K = 32
    gs_pm = torch.randn(B, M, C2, generator=g).to(cuda_dev)
    gs_pm = torch.where(torch.rand(B, M, C2, generator=g).to(cuda_dev) < 0.3, torch.zeros_like(gs_pm), gs_pm)
    ksel_pm = torch.randint(0, K, (B, M, C2), generator=g, dtype=torch.uint8).to(cuda_dev)
    kc2, c1pad = (C2 + 31) // 32, (C1 + 31) // 32 * 32
    wb = torch.zeros((C1, kc2 * 32 + (c1pad if training else 0)), device=cuda_dev)
    wb[:, :C2] = w2.t()
    vvec = None
    if training:
        b2 = (torch.randn(C2, generator=g) * 0.05).to(cuda_dev)
        wb[:, kc2 * 32:kc2 * 32 + C1] = w2.t() @ (b2.view(C2, 1) * w2)
        vvec = (torch.randn(C1, generator=g) * 0.1).to(cuda_dev)
    mean1 = (torch.randn(C1, generator=g) * 0.2).to(cuda_dev)
    rstd1 = (torch.rand(C1, generator=g) + 0.5).to(cuda_dev)
    outs = []
    for fn in (lambda *a: _cabi.call("apb_sa2_bwd_rows_tf32", list(a), cuda_dev), emul.apb_sa2_bwd_rows_tf32):
        du = torch.zeros((B, N, C1), device=cuda_dev)
        sums1 = torch.zeros((C1, 2), dtype=torch.float64, device=cuda_dev)
        sdzdp = torch.zeros((C1, 3), dtype=torch.float64, device=cuda_dev)
        sh1 = torch.zeros(C1, dtype=torch.float64, device=cuda_dev)
        dxyz = None if training else torch.zeros_like(p)
        dq = None if training else torch.zeros_like(q)
        fn(u, p, q, idx, wp, s1, t1, 1, gs_pm, ksel_pm, wb, wb.shape[1], kc2, vvec, mean1, rstd1, du, sums1, sdzdp, sh1, dxyz,
           dq, B, N, M, K, C1, C2)
        outs.append((du, sums1[:, 0], sums1[:, 1], sdzdp, sh1, dxyz, dq))
    torch.cuda.synchronize()
    names = ("du", "sum dz", "sum dz xhat", "sum dz dp", "sum h1", "dxyz", "dq")
    for name, a, b in zip(names, *outs):
        if a is None:
            continue
        assert _rel(a, b) < 5e-3, name
        assert _cos(a, b) > 0.9999, name


@pytest.mark.parametrize("with_gram", [1, 0])
@pytest.mark.parametrize("B,N,M,C1,C2", SA2_SHAPES)
def test_sa2_wgrad_kernel_vs_restatement(cuda_dev, B, N, M, C1, C2, with_gram):
    import _sa2_emul as emul
    from adaptpoint_b200 import _cabi
    g, p, q, idx, u, wp, s1, t1, w2 = _sa2_inputs(cuda_dev, B, N, M, C1, C2, C1 + C2)
    K = 32
    gs_pm = torch.randn(B, M, C2, generator=g).to(cuda_dev)
    ksel_pm = torch.randint(0, K, (B, M, C2), generator=g, dtype=torch.uint8).to(cuda_dev)
    c1pad = (C1 + 31) // 32 * 32
    outs = []
    for fn in (lambda *a: _cabi.call("apb_sa2_wgrad_tf32", list(a), cuda_dev), emul.apb_sa2_wgrad_tf32):
        d = torch.zeros((C2 + (C1 if with_gram else 0), c1pad), device=cuda_dev)
        fn(u, p, q, idx, wp, s1, t1, 1, gs_pm, ksel_pm, d, with_gram, B, N, M, K, C1, C2)
        outs.append(d)
    torch.cuda.synchronize()
    got, want = outs
    assert _rel(got[:C2], want[:C2]) < 5e-3, "S"
    if with_gram:
        assert _rel(got[C2:], want[C2:]) < 5e-3, "H"
    assert torch.count_nonzero(got[:, C1:]).item() == 0, "padding columns stay zero"


def _compare_tf32_two(got, want, pos_grad):
    """Two TF32 GEMMs deep: outputs at 5e-3, gradients by direction and norm (max-over-K routes flip under TF32 noise)."""
    assert _rel(got[0], want[0]) < 5e-3
    assert _cos(got[1], want[1]) > 0.985 and abs(got[1].norm().item() / want[1].norm().item() - 1) < 3e-2, "df"
    if pos_grad:
        assert _cos(got[2], want[2]) > 0.985, "dpos"
    for n, g in want[3].items():
        assert _cos(got[3][n], g) > 0.985, n
        assert abs(got[3][n].norm().item() / g.norm().item() - 1) < 3e-2, n
    for n, b in want[4].items():
        if b.dtype.is_floating_point:
            assert _rel(got[4][n], b) < 5e-3, n


@pytest.mark.parametrize("cin,cout,B,N,training,use_res", [(32, 64, 4, 1024, True, True), (64, 128, 4, 512, True, False),
                                                          (128, 256, 2, 256, True, True), (32, 64, 4, 1024, False, True)])
def test_set_abstraction_two_layer_fused_train_vs_unfused(cuda_dev, cin, cout, B, N, training, use_res):
    """Stride-2 SetAbstraction with sa_layers = 2 (PointNeXt-S): fused (csrc/sa_train2.cu) against the unfused op sequence
    on the strict fp32 torch backend, training (batch statistics) and eval mode (running statistics + coordinate grads)."""
    from adaptpoint_b200 import fused_train
    from adaptpoint_b200.pointnext import SetAbstraction
    torch.manual_seed(cin + N)
    sa = SetAbstraction(cin, cout, layers=2, stride=2, use_res=use_res,
                        group_args={"NAME": "ballquery", "radius": 0.25, "nsample": 32, "normalize_dp": True},
                        norm_args={"norm": "bn"}, act_args={"act": "relu"}).to(cuda_dev)
    with torch.no_grad():
        for blk in sa.convs:
            bn = blk[1]
            c = bn.weight.numel()
            bn.weight.copy_(torch.randn(c) * 0.7 + 0.3)
            bn.bias.copy_(torch.randn(c) * 0.2)
            bn.running_mean.copy_(torch.randn(c) * 0.1)
            bn.running_var.copy_(torch.rand(c) + 0.5)
    sa.train(training)
    p, f = _inputs(cuda_dev, B, N, cin, 13)
    ref = copy.deepcopy(sa)
    want = _run(ref, p, f, "torch", False, 4, pos_grad=not training)
    launched = []
    from adaptpoint_b200 import _cabi
    real = _cabi.call
    _cabi.call = lambda name, a, d: (launched.append(name), real(name, a, d))[1]
    try:
        got = _run(sa, p, f, "tc", True, 4, pos_grad=not training)
    finally:
        _cabi.call = real
    assert "apb_sa2_fwd_tf32" in launched and "apb_sa2_bwd_rows_tf32" in launched and "apb_sa2_wgrad_tf32" in launched
    assert "apb_group_f32" not in launched, "the grouped tensor must not be built"
    _compare_tf32_two(got, want, not training)


@pytest.mark.parametrize("B,N,M,C1", [(3, 256, 5, 32), (4, 512, 128, 64), (2, 130, 33, 20), (8, 1024, 512, 128)])
def test_sa_ustats_closed_form_equals_gathered_statistics(cuda_dev, B, N, M, C1):
    """apb_sa_ustats_f32 (sum y, sum y^2 from cnt / sdp / moments, one pass over U) against apb_group_ystats_f32 (the same
    sums taken over the B*M*K gathered rows) and the float64 restatement."""
    import _sa2_emul as emul
    from adaptpoint_b200 import _cabi
    _, p, q, idx, u, wp, _, _, _ = _sa2_inputs(cuda_dev, B, N, M, C1, 8, C1 + N)
    K = 32
    cnt = torch.zeros((B, N), device=cuda_dev)
    sdp = torch.zeros((B, N, 3), device=cuda_dev)
    mom = torch.zeros(12, dtype=torch.float64, device=cuda_dev)
    _cabi.call("apb_sa1_bwd_count_f32", [p, q, idx, cnt, sdp, mom, B, N, M, K], cuda_dev)
    got = torch.zeros((C1, 2), dtype=torch.float64, device=cuda_dev)
    _cabi.call("apb_sa_ustats_f32", [u, cnt, sdp, wp, mom, got, B, N, C1], cuda_dev)
    direct = torch.zeros((C1, 2), dtype=torch.float64, device=cuda_dev)
    _cabi.call("apb_group_ystats_f32", [u, p, q, idx, wp, None, None, None, direct, B, N, M, K, C1], cuda_dev)
    want = torch.zeros((C1, 2), dtype=torch.float64, device=cuda_dev)
    emul.apb_group_ystats_f32(u, p, q, idx, wp, None, None, None, want, B, N, M, K, C1)
    torch.cuda.synchronize()
    n = B * M * K
    # mean and variance, the quantities BatchNorm uses
    for name, s in (("closed form", got), ("gathered", direct)):
        mean, ex2 = s[:, 0] / n, s[:, 1] / n
        wmean, wex2 = want[:, 0] / n, want[:, 1] / n
        assert (mean - wmean).abs().max().item() < 1e-5 * (1 + wmean.abs().max().item()), name
        assert ((ex2 - mean * mean) - (wex2 - wmean * wmean)).abs().max().item() < 1e-4 * (wex2 - wmean * wmean).max().item(), name
