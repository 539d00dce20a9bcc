"""apb_group_minmax_f32 / ops.group_minmax (grouping fused with max and min over the neighbours) against the unfused
composition grouping_operation -> torch max/min, and PointsetGrouper's fused normalise-then-max against the
reference formulation (generator_component4_15.py:368-431).  Values are bit-exact; gradients 1e-5 relative."""
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("torch_backend_default")]


@pytest.mark.parametrize("B,C,N,M,K", [(4, 64, 2048, 1024, 24), (2, 5, 1000, 333, 7), (3, 128, 512, 128, 32),
                                       (1, 1, 40, 40, 3), (2, 2, 20000, 50, 16), (1, 3, 40000, 16, 4)])
def test_group_minmax_matches_unfused(cuda_dev, B, C, N, M, K):
    from adaptpoint_b200 import ops
    gen = torch.Generator().manual_seed(N + K)
    f = torch.randn(B, C, N, generator=gen).to(cuda_dev).requires_grad_(True)
    idx = torch.randint(0, N, (B, M, K), generator=gen, dtype=torch.int32).to(cuda_dev)
    idx[:, :, K // 2:] = idx[:, :, :1]                       # ball-query style padding with the first neighbour
    gmx = torch.randn(B, C, M, generator=gen).to(cuda_dev)
    gmn = torch.randn(B, C, M, generator=gen).to(cuda_dev)
    g = ops.grouping_operation(f, idx)
    want_mx, want_mn = g.max(dim=3)[0], g.min(dim=3)[0]
    (want_mx * gmx + want_mn * gmn).sum().backward()
    want_grad = f.grad.clone()
    f.grad = None
    mx, mn = ops.group_minmax(f, idx)
    assert torch.equal(mx, want_mx) and torch.equal(mn, want_mn)
    (mx * gmx + mn * gmn).sum().backward()
    assert (f.grad - want_grad).abs().max() <= 1e-5 * want_grad.abs().max().clamp_min(1.0)


def test_pointset_grouper_fused_vs_reference_formula(cuda_dev):
    from adaptpoint_b200 import ops
    from adaptpoint_b200.imitator import PointsetGrouper
    from adaptpoint_b200.synth import make_clouds
    torch.manual_seed(0)
    B, C, N = 4, 32, 1024
    xyz = make_clouds(B, N, "shell", 5).to(cuda_dev)
    feats = torch.randn(B, C, N, device=cuda_dev, requires_grad=True)
    m = PointsetGrouper(C, 2, 24, 0.2).to(cuda_dev)
    with torch.no_grad():
        m.affine_alpha.copy_(torch.randn_like(m.affine_alpha))      # both signs: exercises the min branch
        m.affine_beta.copy_(torch.randn_like(m.affine_beta))
    new_xyz, out = m(xyz, feats)
    gy = torch.randn_like(out)
    out.backward(gy)
    got = [feats.grad.clone(), m.affine_alpha.grad.clone(), m.affine_beta.grad.clone()]
    feats.grad = None
    m.zero_grad()
    # reference formulation on the materialised (B,C,np,k) tensor
    fps_idx = ops.furthest_point_sample(xyz, N // 2).long()
    q = torch.gather(xyz, 1, fps_idx.unsqueeze(-1).expand(-1, -1, 3))
    centre = torch.gather(feats, 2, fps_idx.unsqueeze(1).expand(-1, C, -1))
    idx = ops.ball_query(0.2, 24, xyz, q)
    grouped = ops.grouping_operation(feats, idx)
    want = (m.affine_alpha.view(1, C, 1, 1) * (grouped - centre.unsqueeze(-1)) + m.affine_beta.view(1, C, 1, 1)).max(dim=3)[0]
    want.backward(gy)
    assert torch.equal(new_xyz, q)
    assert torch.equal(out, want)
    for a, b in zip(got, [feats.grad, m.affine_alpha.grad, m.affine_beta.grad]):
        assert (a - b).abs().max() <= 1e-4 * b.abs().max().clamp_min(1.0)
