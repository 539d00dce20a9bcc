"""Host-side algebra of the fused two-conv set abstraction (adaptpoint_b200/fused_train.py: SaTwoLayerFn) on CPU: every
C-ABI call is swapped for its torch restatement (tests/_sa2_emul.py), and forward value + every gradient are compared
with plain torch autograd of the reference's op chain (pointnext.py:164-166, conv.py:24-60).  The CUDA kernels themselves
are compared with the same restatements in tests/test_fused_train_gpu.py."""
import pytest
import torch

from adaptpoint_b200 import _cabi, fused_train
import _sa2_emul as emul


def _case(seed, B=2, N=48, M=12, K=32, cin=8, C1=16, C2=24):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    p = r(B, N, 3) * 0.3
    q = p[:, :M].clone()
    f = r(B, cin, N)
    idx = torch.randint(0, N, (B, M, K), generator=g, dtype=torch.int32)
    w1 = r(C1, cin + 3, 1, 1) * 0.4
    w2 = r(C2, C1, 1, 1) * 0.4
    bn = lambda c: (1 + 0.3 * r(c), 0.2 * r(c), 0.1 * r(c), 0.5 + torch.rand(c, generator=g))
    return p, q, f, idx, w1, bn(C1), w2, bn(C2)


@pytest.mark.parametrize("training,relu2,coords", [(True, True, False), (True, False, False), (False, True, True),
                                                    (False, False, True)])
def test_two_conv_algebra_matches_autograd(monkeypatch, training, relu2, coords):
    monkeypatch.setattr(_cabi, "call", emul.call)
    monkeypatch.setattr(fused_train, "EXACT_GEMM", True)
    p, q, f, idx, w1, (g1, b1, rm1, rv1), w2, (g2, b2, rm2, rv2) = _case(3)
    # negative BatchNorm-2 weights exercise the min branch of the monotone max
    g2 = g2 * torch.where(torch.arange(g2.numel()) % 3 == 0, -1.0, 1.0)
    inv_r = 1.0 / 0.35
    leaves = [t.clone().requires_grad_(True) for t in (f, w1, g1, b1, w2, g2, b2)]
    pp, qq = p.clone().requires_grad_(coords), q.clone().requires_grad_(coords)
    rms = [t.clone() for t in (rm1, rv1, rm2, rv2)]
    out = fused_train.SaTwoLayerFn.apply(pp, qq, leaves[0], idx, leaves[1], leaves[2], leaves[3], rms[0], rms[1], 0.1, 1e-5,
                                         leaves[4], leaves[5], leaves[6], rms[2], rms[3], 0.1, 1e-5, relu2, training, None,
                                         inv_r)
    go = torch.randn(out.shape, generator=torch.Generator().manual_seed(9))
    out.backward(go)

    ref_leaves = [t.double().clone().requires_grad_(True) for t in (f, w1, g1, b1, w2, g2, b2)]
    rp, rq = p.double().clone().requires_grad_(coords), q.double().clone().requires_grad_(coords)
    rrm = [t.double().clone() for t in (rm1, rv1, rm2, rv2)]
    ref = emul.reference_block(rp, rq, ref_leaves[0], idx, ref_leaves[1], ref_leaves[2], ref_leaves[3], ref_leaves[4],
                               ref_leaves[5], ref_leaves[6], relu2, training, inv_r, rrm[0], rrm[1], rrm[2], rrm[3])
    ref.backward(go.double())

    def close(name, a, b, tol=2e-4):
        a, b = a.double(), b.double()
        err = (a - b).abs().max().item()
        assert err <= tol * (1.0 + b.abs().max().item()), f"{name}: max |diff| {err:.3e} vs scale {b.abs().max().item():.3e}"

    close("out", out, ref)
    for name, a, b in zip(("f", "w1", "g1", "b1", "w2", "g2", "b2"), leaves, ref_leaves):
        assert a.grad is not None, name
        close("grad " + name, a.grad, b.grad)
    if coords:
        close("grad p", pp.grad, rp.grad)
        close("grad q", qq.grad, rq.grad)
    if training:   # running statistics updated like torch's
        for name, a, b in zip(("rm1", "rv1", "rm2", "rv2"), rms, rrm):
            close(name, a, b)
