"""Training-path 1x1 convolutions on the tensor cores (tcconv.PointwiseConvFn: apb_pwconv_fwd_bf16 for y and gx,
apb_pwconv_wgrad_tf32 for gW) against a float64 contraction; tolerance 3e-3 (TF32 operands) / 2e-2 (bf16 operands)
relative, fp32 accumulation."""
import pytest
import torch
import torch.nn as nn

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("torch_backend_default")]
TOL = 2e-2


def _rel(got, want):
    return ((got - want).abs().max() / want.abs().max().clamp_min(1e-6)).item()


@pytest.mark.parametrize("B,cin,cout,shape,bias", [
    (4, 35, 32, (512, 32), False),     # SA stage-1 conv1 on the grouped tensor
    (4, 32, 64, (512, 32), False),     # conv2
    (2, 259, 256, (64, 32), False),    # stage 4
    (2, 256, 512, (64, 32), True),
    (3, 4, 32, (1024,), True),         # head Conv1d
    (2, 128, 1024, (2048,), True),     # discriminator: Cout > 512 is split
    (2, 515, 64, (96,), False),        # Cin > 512: data gradient is split
    (2, 64, 3, (1000,), False),        # tiny Cout
    (2, 16, 24, (333,), True),         # P % 4 != 0: weight gradient falls back to einsum
])
@pytest.mark.parametrize("prec,tol", [("tf32", 3e-3), ("bf16", TOL)])
def test_pointwise_conv_fwd_bwd(cuda_dev, B, cin, cout, shape, bias, prec, tol):
    from adaptpoint_b200 import tcconv
    tcconv.set_conv_precision(prec)
    gen = torch.Generator().manual_seed(cin + cout)
    Conv = tcconv.TcConv2d if len(shape) == 2 else tcconv.TcConv1d
    conv = Conv(cin, cout, 1, bias=bias).to(cuda_dev)
    x = torch.randn((B, cin) + shape, generator=gen).to(cuda_dev).requires_grad_(True)
    gy = torch.randn((B, cout) + shape, generator=gen).to(cuda_dev)
    # comparand: float64 contraction (not cuDNN, whose own TF32 switch would make the reference order-dependent)
    xd, gd = x.detach().double().flatten(2), gy.double().flatten(2)
    wd = conv.weight.detach().double().reshape(cout, cin)
    want = torch.einsum("oi,bip->bop", wd, xd)
    if bias:
        want = want + conv.bias.detach().double().view(1, cout, 1)
    want = want.view((B, cout) + shape).float()
    ref = [torch.einsum("oi,bop->bip", wd, gd).view((B, cin) + shape).float(),
           torch.einsum("bop,bip->oi", gd, xd).view(conv.weight.shape).float(),
           None if not bias else gd.sum(dim=(0, 2)).float()]
    with tcconv.conv_backend("tc"):
        got = conv(x)
        got.backward(gy)
    tcconv.set_conv_precision("tf32")
    assert got.shape == want.shape
    assert _rel(got, want) < tol
    assert _rel(x.grad, ref[0]) < tol
    assert _rel(conv.weight.grad, ref[1]) < 5e-3       # wgrad is tf32 in both modes
    if bias:
        assert _rel(conv.bias.grad, ref[2]) < 1e-4


def test_wgrad_large_k_split(cuda_dev):
    """Many K chunks per CTA and two Cin tiles."""
    from adaptpoint_b200 import _cabi
    gen = torch.Generator().manual_seed(0)
    B, cin, cout, P = 8, 300, 200, 4096
    x = torch.randn(B, cin, P, generator=gen).to(cuda_dev)
    gy = torch.randn(B, cout, P, generator=gen).to(cuda_dev)
    gw = torch.zeros(cout, cin, device=cuda_dev)
    _cabi.call("apb_pwconv_wgrad_tf32", [gy, x, gw, B, cin, cout, P], cuda_dev)
    want = torch.einsum("bop,bip->oi", gy.double(), x.double()).float()
    assert _rel(gw, want) < 5e-3


def test_model_step_tc_vs_fp32(cuda_dev):
    """PointNeXt-S training-mode loss and a few gradients: tensor-core convs vs cuDNN fp32, same parameters."""
    from adaptpoint_b200 import tcconv
    from adaptpoint_b200.models import build_model
    from adaptpoint_b200.synth import make_clouds
    torch.manual_seed(0)
    model = build_model("pointnext-s").to(cuda_dev).train()
    xyz = make_clouds(32, 1024, "shell", 3).to(cuda_dev)
    x = torch.cat([xyz, xyz[:, :, 1:2] - xyz[:, :, 1:2].amin(1, keepdim=True)], -1).transpose(1, 2).contiguous()
    y = torch.arange(32, device=cuda_dev) % 15
    outs = {}
    for backend in ("torch", "tc"):
        model.zero_grad()
        torch.manual_seed(7)                       # same Dropout masks in both runs
        with tcconv.conv_backend(backend):
            logits, loss = model.get_logits_loss({"pos": xyz, "x": x}, y)
            loss.backward()
        grads = {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None}
        outs[backend] = (logits.detach().clone(), loss.item(), grads)
    assert abs(outs["tc"][1] - outs["torch"][1]) < 2e-3 * max(1.0, abs(outs["torch"][1]))
    assert _rel(outs["tc"][0], outs["torch"][0]) < 5e-2   # head BatchNorm1d over 8 samples amplifies ~10x
    # weight gradients through 12 stacked layers, training-mode BatchNorm, max-pool routing: tf32 operands keep the
    # direction (bf16 operands do not: cos ~0.75, see tcconv.set_conv_precision)
    for n, g in outs["torch"][2].items():
        gt = outs["tc"][2][n]
        if g.numel() < 64 or g.norm() < 1e-6:
            continue
        cos = torch.nn.functional.cosine_similarity(g.flatten(), gt.flatten(), dim=0).item()
        assert cos > 0.9, (n, cos)   # earliest layers see the compounded routing noise (0.96 measured); later ones > 0.99


@pytest.mark.parametrize("B,cin,cout,P,bias", [(4, 128, 1024, 2048, True), (3, 64, 96, 777, False), (2, 3, 64, 100, True)])
def test_conv_max_fused_vs_torch(cuda_dev, B, cin, cout, P, bias):
    """tcconv.conv_max (GEMM slices -> apb_rowmax_f32, backward through the arg-max columns only) against
    relu(conv) -> max over the points through torch autograd (point_discriminator.py:166-191)."""
    from adaptpoint_b200 import tcconv
    gen = torch.Generator().manual_seed(P)
    conv = tcconv.TcConv2d(cin, cout, 1, bias=bias).to(cuda_dev)
    x = torch.randn(B, cin, 1, P, generator=gen).to(cuda_dev).requires_grad_(True)
    gy = torch.randn(B, cout, generator=gen).to(cuda_dev)
    with tcconv.conv_backend("torch"):
        want = torch.relu(conv(x)).flatten(2).max(dim=2)[0]
        want.backward(gy)
    ref = [x.grad.clone(), conv.weight.grad.clone(), None if not bias else conv.bias.grad.clone()]
    x.grad = None
    conv.zero_grad()
    with tcconv.conv_backend("tc"):
        got = torch.relu(tcconv.conv_max(conv, x))
        got.backward(gy)
    assert got.shape == want.shape
    assert _rel(got, want) < 3e-3
    # tf32 rounding can move an arg-max between near-tied columns: compare gradients in aggregate
    cos = torch.nn.functional.cosine_similarity
    assert cos(x.grad.flatten(), ref[0].flatten(), dim=0) > 0.98
    assert cos(conv.weight.grad.flatten(), ref[1].flatten(), dim=0) > 0.98
    if bias:
        assert _rel(conv.bias.grad, ref[2]) < 1e-4


def test_rowmax_exact(cuda_dev):
    """apb_rowmax_f32: bit-exact max and FIRST arg-max per row, incl. ties, odd P and unaligned rows."""
    from adaptpoint_b200 import _cabi
    gen = torch.Generator().manual_seed(3)
    for rows, P in ((64, 2048), (37, 777), (5, 3), (1, 1)):
        y = torch.randint(-8, 8, (rows, P), generator=gen).float().to(cuda_dev)     # many exact ties
        v = torch.empty(rows, device=cuda_dev)
        a = torch.empty(rows, dtype=torch.int32, device=cuda_dev)
        _cabi.call("apb_rowmax_f32", [y, v, a, rows, P], cuda_dev)
        want_v = y.max(dim=1)[0]
        first = (y == want_v[:, None]).float().argmax(dim=1)
        assert torch.equal(v, want_v)
        assert torch.equal(a.long(), first)


@pytest.mark.parametrize("B,cin,cout,P", [(32, 128, 512, 2048), (4, 35, 32, 16384), (2, 3, 64, 512), (3, 259, 256, 2048),
                                          (2, 64, 1024, 1000), (5, 96, 40, 132), (1, 515, 72, 4), (2, 16, 272, 8196)])
def test_tma_gemm_store_stats_pointmajor(cuda_dev, B, cin, cout, P):
    """csrc/tc_cm_gemm.cu (both operands by TMA, MN-major activation tile): channel-major store + fused BatchNorm
    statistics, and the point-major store, against a float64 contraction (TF32 operands: 3e-3)."""
    from adaptpoint_b200 import _cabi, tcconv
    gen = torch.Generator().manual_seed(cin * 7 + cout)
    x = torch.randn(B, cin, P, generator=gen).to(cuda_dev)
    w = (torch.randn(cout, cin, generator=gen) / cin ** 0.5).to(cuda_dev)
    bias = torch.randn(cout, generator=gen).to(cuda_dev)
    wb, kp = tcconv.pack_f32(w)
    want = torch.einsum("oi,bip->bop", w.double(), x.double()) + bias.double().view(1, cout, 1)
    y = torch.empty(B, cout, P, device=cuda_dev)
    stats = torch.zeros(cout, 2, dtype=torch.float64, device=cuda_dev)
    _cabi.call("apb_pwconv_ex_tf32", [x, wb, None, bias, 1, y, 0, stats, None, B, cin, cout, P, kp], cuda_dev)
    assert _rel(y, want.clamp_min(0).float()) < 3e-3
    n = B * P
    mean_w, sq_w = want.sum(dim=(0, 2)) / n, (want * want).sum(dim=(0, 2)) / n
    assert ((stats[:, 0] / n - mean_w).abs().max() / mean_w.abs().max()).item() < 2e-3      # statistics are pre-ReLU
    assert ((stats[:, 1] / n - sq_w).abs().max() / sq_w.abs().max()).item() < 2e-3
    ypm = torch.empty(B * P, cout, device=cuda_dev)
    _cabi.call("apb_pwconv_ex_tf32", [x, wb, None, None, 0, ypm, 1, None, None, B, cin, cout, P, kp], cuda_dev)
    want_pm = (want - bias.double().view(1, cout, 1)).permute(0, 2, 1).reshape(B * P, cout).float()
    assert _rel(ypm, want_pm) < 3e-3


@pytest.mark.parametrize("B,cin,cout,P", [(32, 128, 1024, 2048), (3, 64, 96, 1000), (2, 12, 64, 100), (4, 128, 300, 388)])
def test_tma_gemm_max_over_points_exact(cuda_dev, B, cin, cout, P):
    """The max / first-arg-max epilogue must agree EXACTLY with a max over the activation the store epilogue of the same
    kernel writes (same MMA order => same bits)."""
    from adaptpoint_b200 import _cabi, tcconv
    gen = torch.Generator().manual_seed(P + cout)
    x = torch.randn(B, cin, P, generator=gen).to(cuda_dev)
    x[:, :, P // 2:] = x[:, :, :P - P // 2]                      # duplicated points: exact ties, first index must win
    w = (torch.randn(cout, cin, generator=gen) / cin ** 0.5).to(cuda_dev)
    bias = torch.randn(cout, generator=gen).to(cuda_dev)
    wb, kp = tcconv.pack_f32(w)
    y = torch.empty(B, cout, P, device=cuda_dev)
    _cabi.call("apb_pwconv_fwd_tf32", [x, wb, None, bias, 0, y, B, cin, cout, P, kp], cuda_dev)
    packed = torch.zeros(B, cout, dtype=torch.int64, device=cuda_dev)
    _cabi.call("apb_pwconv_ex_tf32", [x, wb, None, bias, 0, None, 0, None, packed, B, cin, cout, P, kp], cuda_dev)
    val = torch.empty(B, cout, device=cuda_dev)
    arg = torch.empty(B, cout, dtype=torch.int32, device=cuda_dev)
    _cabi.call("apb_pmax_decode", [packed, val, arg, B * cout], cuda_dev)
    want_v = y.max(dim=2)[0]
    first = (y == want_v[:, :, None]).float().argmax(dim=2)
    assert torch.equal(val, want_v)
    assert torch.equal(arg.long(), first)


def test_conv_bn_block_fused_stats_matches_unfused(cuda_dev):
    """ConvBlock (Conv -> BatchNorm -> ReLU) with the statistics taken in the GEMM epilogue vs the separate statistics
    pass (apb_bn_stats_f32) over the same conv output: forward, gradients and running statistics."""
    import copy
    from adaptpoint_b200 import layers, norm, tcconv
    torch.manual_seed(0)
    blk = layers.create_convblock2d(35, 64, norm_args={"norm": "bn"}, act_args={"act": "relu"}).to(cuda_dev).train()
    ref = copy.deepcopy(blk)
    x = torch.randn(4, 35, 256, 32, device=cuda_dev)
    gy = torch.randn(4, 64, 256, 32, device=cuda_dev)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    with tcconv.conv_backend("tc"):
        ya = blk(xa)                                              # fused statistics
        ya.backward(gy)
        yb = ref[1].forward_act(ref[0](xb), True)                 # conv, then BatchNorm with its own statistics pass
        yb.backward(gy)
    assert _rel(ya, yb) < 1e-4
    assert _rel(xa.grad, xb.grad) < 1e-4
    assert _rel(blk[0].weight.grad, ref[0].weight.grad) < 1e-4
    assert _rel(blk[1].weight.grad, ref[1].weight.grad) < 1e-4
    assert _rel(blk[1].running_var, ref[1].running_var) < 1e-4 and _rel(blk[1].running_mean, ref[1].running_mean) < 1e-4
