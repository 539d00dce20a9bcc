"""BatchNorm(+ReLU) and BatchNorm->max kernels (csrc/norm.cu, adaptpoint_b200/norm.py) against torch's
nn.BatchNorm / ReLU / max + autograd, fp32, tolerance 1e-4 relative (different summation order only)."""
import copy

import pytest
import torch
import torch.nn as nn

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("torch_backend_default")]
TOL = 2e-4


def _rel(got, want):
    return ((got - want).abs().max() / want.abs().max().clamp_min(1e-6)).item()


def _mk_bn(C, dim, gen, dev):
    from adaptpoint_b200 import norm
    bn = (norm.TcBatchNorm2d if dim == 2 else norm.TcBatchNorm1d)(C)
    bn.weight.data.copy_(torch.randn(C, generator=gen))        # negative gammas included
    bn.bias.data.copy_(torch.randn(C, generator=gen) * 0.3)
    bn.running_mean.copy_(torch.randn(C, generator=gen) * 0.1)
    bn.running_var.copy_(torch.rand(C, generator=gen) + 0.5)
    return bn.to(dev)


@pytest.mark.parametrize("shape,relu", [((4, 32, 512, 32), True), ((3, 67, 100, 32), False), ((2, 16, 1023), True),
                                        ((8, 512, 64), True), ((5, 9, 7, 3), True), ((2, 64, 20000), False),
                                        ((32, 1024, 4), True), ((4, 300, 1), False), ((3, 40, 65), True)])
def test_bn_act_matches_torch(cuda_dev, shape, relu):
    from adaptpoint_b200 import tcconv
    gen = torch.Generator().manual_seed(sum(shape))
    dim = 2 if len(shape) == 4 else 1
    bn = _mk_bn(shape[1], dim, gen, cuda_dev).train()
    ref = copy.deepcopy(bn)
    x = (torch.randn(shape, generator=gen) * 2 + 0.5).to(cuda_dev)
    gy = torch.randn(shape, generator=gen).to(cuda_dev)
    xr = x.clone().requires_grad_(True)
    want = ref(xr)
    if relu:
        want = torch.relu(want)
    want.backward(gy)
    xt = x.clone().requires_grad_(True)
    with tcconv.conv_backend("tc"):
        got = bn.forward_act(xt, relu)
        got.backward(gy)
    assert _rel(got, want) < TOL
    assert _rel(xt.grad, xr.grad) < TOL
    assert _rel(bn.weight.grad, ref.weight.grad) < TOL
    assert _rel(bn.bias.grad, ref.bias.grad) < TOL
    assert _rel(bn.running_mean, ref.running_mean) < TOL
    assert _rel(bn.running_var, ref.running_var) < TOL
    assert int(bn.num_batches_tracked) == int(ref.num_batches_tracked) == 1


@pytest.mark.parametrize("shape,relu,training", [((4, 64, 256, 32), False, True), ((2, 35, 100, 32), True, True),
                                                 ((3, 512, 1, 64), False, True), ((2, 16, 50, 24), True, True),
                                                 ((2, 7, 33, 5), True, True), ((4, 64, 256, 32), False, False),
                                                 ((2, 32, 128, 32), True, False)])
def test_bn_max_matches_torch(cuda_dev, shape, relu, training):
    from adaptpoint_b200 import norm, tcconv
    gen = torch.Generator().manual_seed(sum(shape) + relu)
    bn = _mk_bn(shape[1], 2, gen, cuda_dev).train(training)
    ref = copy.deepcopy(bn)
    y = (torch.randn(shape, generator=gen) * 1.5 - 0.2).to(cuda_dev)
    g = torch.randn(shape[:3], generator=gen).to(cuda_dev)
    yr = y.clone().requires_grad_(True)
    z = ref(yr)
    if relu:
        z = torch.relu(z)
    want = z.max(dim=-1)[0]
    want.backward(g)
    yt = y.clone().requires_grad_(True)
    with tcconv.conv_backend("tc"):
        got = norm.bn_max(yt, bn, relu)
        got.backward(g)
    assert got.shape == want.shape
    assert _rel(got, want) < TOL
    assert _rel(yt.grad, yr.grad) < 5e-4
    assert _rel(bn.weight.grad, ref.weight.grad) < 5e-4
    assert _rel(bn.bias.grad, ref.bias.grad) < 5e-4
    assert _rel(bn.running_mean, ref.running_mean) < TOL
    assert _rel(bn.running_var, ref.running_var) < TOL


def test_model_grads_tc_norm_vs_torch(cuda_dev):
    """PointNeXt-S training step: our BatchNorm / bn_max kernels (fp32) with cuDNN convs vs the stock torch path."""
    from adaptpoint_b200 import norm, tcconv
    from adaptpoint_b200.models import build_model
    from adaptpoint_b200.synth import make_clouds
    torch.manual_seed(0)
    model = build_model("pointnext-s").to(cuda_dev).train()
    xyz = make_clouds(32, 1024, "shell", 3).to(cuda_dev)
    x = torch.cat([xyz, xyz[:, :, 1:2] - xyz[:, :, 1:2].amin(1, keepdim=True)], -1).transpose(1, 2).contiguous()
    y = torch.arange(32, device=cuda_dev) % 15
    state = copy.deepcopy(model.state_dict())
    outs = {}
    for mode in ("torch", "tcnorm"):
        model.load_state_dict(state)
        model.zero_grad()
        torch.manual_seed(7)
        if mode == "tcnorm":
            # tc backend for the norm kernels, convs forced back to cuDNN fp32 so only the norm path differs
            # (the fused set-abstraction path runs its own TF32 GEMMs: off, it has its own tests)
            from adaptpoint_b200 import fused_train
            orig, orig_fused = tcconv._use_tc, fused_train.ENABLED
            tcconv._use_tc = lambda m, t: False
            fused_train.ENABLED = False
            try:
                with tcconv.conv_backend("tc"):
                    logits, loss = model.get_logits_loss({"pos": xyz, "x": x}, y)
                    loss.backward()
            finally:
                tcconv._use_tc = orig
                fused_train.ENABLED = orig_fused
        else:
            logits, loss = model.get_logits_loss({"pos": xyz, "x": x}, y)
            loss.backward()
        outs[mode] = (logits.detach().clone(), {n: p.grad.clone() for n, p in model.named_parameters()
                                                if p.grad is not None},
                      {n: b.clone() for n, b in model.named_buffers()})
    assert _rel(outs["tcnorm"][0], outs["torch"][0]) < 2e-2   # B=8: the head BatchNorm1d amplifies reordering noise
    for n, gref in outs["torch"][1].items():
        if gref.norm() < 1e-6:
            continue
        cos = torch.nn.functional.cosine_similarity(gref.flatten(), outs["tcnorm"][1][n].flatten(), dim=0).item()
        assert cos > 0.99, (n, cos)
    for n, bref in outs["torch"][2].items():
        assert _rel(outs["tcnorm"][2][n].float(), bref.float()) < 2e-2, n


@pytest.mark.parametrize("shape,relu,param_grads", [((4, 32, 512, 32), True, True), ((3, 67, 1001), False, True),
                                                    ((2, 64, 128, 32), True, False), ((32, 512, 1), True, True),
                                                    ((4, 40, 7), False, True)])
def test_bn_eval_autograd(cuda_dev, shape, relu, param_grads):
    """Eval-mode BatchNorm (+ReLU) under autograd (the feedback-loss path), with and without parameter gradients."""
    from adaptpoint_b200 import tcconv
    gen = torch.Generator().manual_seed(sum(shape))
    bn = _mk_bn(shape[1], 2 if len(shape) == 4 else 1, gen, cuda_dev).eval()
    ref = copy.deepcopy(bn)
    for m in (bn, ref):
        for p in m.parameters():
            p.requires_grad_(param_grads)
    x = torch.randn(shape, generator=gen).to(cuda_dev)
    gy = torch.randn(shape, generator=gen).to(cuda_dev)
    xr = x.clone().requires_grad_(True)
    want = ref(xr)
    if relu:
        want = torch.relu(want)
    want.backward(gy)
    xt = x.clone().requires_grad_(True)
    with tcconv.conv_backend("tc"):
        got = bn.forward_act(xt, relu)
        got.backward(gy)
    assert _rel(got, want) < TOL and _rel(xt.grad, xr.grad) < TOL
    if param_grads:
        assert _rel(bn.weight.grad, ref.weight.grad) < TOL and _rel(bn.bias.grad, ref.bias.grad) < TOL
    else:
        assert bn.weight.grad is None
