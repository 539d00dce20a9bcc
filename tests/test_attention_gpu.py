"""ops.self_attention (apb_attn_{fwd,bwd}_f32, flash-style fp32, head_dim 16) against the reference formulation
softmax(q k^T / sqrt(d)) v of Anchor_selfattention (generator_component4_15.py:434-480) evaluated in float64.
Tolerance: 1e-5 relative (fp32 path, north_star) on outputs, 5e-5 on gradients (longer fp32 accumulation chains)."""
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("torch_backend_default")]


def _ref(q, k, v, H):
    B, M, C = q.shape
    sp = lambda t: t.reshape(B, M, H, C // H).permute(0, 2, 1, 3)
    a = torch.softmax(sp(q) @ sp(k).transpose(-1, -2) / (C // H) ** 0.5, dim=-1)
    return (a @ sp(v)).permute(0, 2, 1, 3).reshape(B, M, C)


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-12)).item()


@pytest.mark.parametrize("B,M,H,scale", [(2, 2048, 4, 1.0), (3, 4, 4, 1.0), (2, 333, 2, 3.0), (1, 129, 1, 0.3), (2, 1, 4, 1.0)])
def test_self_attention_fwd_bwd(cuda_dev, B, M, H, scale):
    from adaptpoint_b200 import ops
    gen = torch.Generator().manual_seed(M)
    q, k, v = [(scale * torch.randn(B, M, H * 16, generator=gen)).to(cuda_dev).requires_grad_(True) for _ in range(3)]
    go = torch.randn(B, M, H * 16, generator=gen).to(cuda_dev)
    o = ops.self_attention(q, k, v, H)
    o.backward(go)
    got = [o.detach(), q.grad.clone(), k.grad.clone(), v.grad.clone()]
    qd, kd, vd = [t.detach().double().requires_grad_(True) for t in (q, k, v)]
    od = _ref(qd, kd, vd, H)
    od.backward(go.double())
    want = [od.detach(), qd.grad, kd.grad, vd.grad]
    assert _rel(got[0], want[0]) < 1e-5
    for g, w in zip(got[1:], want[1:]):
        assert _rel(g, w) < 5e-5


@pytest.mark.parametrize("B,M,H,D", [(32, 4, 4, 256), (2, 32, 2, 40), (3, 1, 4, 7), (2, 17, 3, 100)])
def test_self_attention_few_tokens_any_head_dim(cuda_dev, B, M, H, D):
    """apb_attn_small_{fwd,bwd}_f32: the imitator's anchor head (4 anchors, head_dim 256) and other widths."""
    from adaptpoint_b200 import ops
    gen = torch.Generator().manual_seed(M * D)
    q, k, v = [torch.randn(B, M, H * D, generator=gen).to(cuda_dev).requires_grad_(True) for _ in range(3)]
    go = torch.randn(B, M, H * D, generator=gen).to(cuda_dev)
    o = ops.self_attention(q, k, v, H)
    o.backward(go)
    got = [o.detach(), q.grad.clone(), k.grad.clone(), v.grad.clone()]
    qd, kd, vd = [t.detach().double().requires_grad_(True) for t in (q, k, v)]
    od = _ref(qd, kd, vd, H)
    od.backward(go.double())
    assert _rel(got[0], od.detach()) < 1e-5
    for g, w in zip(got[1:], [qd.grad, kd.grad, vd.grad]):
        assert _rel(g, w) < 5e-5


def test_self_attention_outside_kernel_family_raises(cuda_dev):
    from adaptpoint_b200 import _cabi, ops
    x = torch.randn(1, 64, 4 * 24, device=cuda_dev)
    with pytest.raises(_cabi.ApbError):
        ops.self_attention(x, x, x, 4)


def test_anchor_selfattention_module_matches_sdpa(cuda_dev):
    """The module on CUDA (our kernel) vs the same module evaluated through torch SDPA on the CPU."""
    from adaptpoint_b200.imitator import Anchor_selfattention
    torch.manual_seed(0)
    m = Anchor_selfattention(64, 4).eval()
    x, xyz = torch.randn(2, 300, 64), torch.randn(2, 300, 3)
    want = m(x, xyz)
    got = m.to(cuda_dev)(x.to(cuda_dev), xyz.to(cuda_dev)).cpu()
    assert _rel(got, want) < 2e-3      # the surrounding 1x1 convs run with tf32 operands on the GPU
