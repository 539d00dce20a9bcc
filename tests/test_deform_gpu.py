"""Fused deformation kernel (csrc/deform.cu, via imitator.deform_mask through the C-ABI) against the torch
restatement of the reference's op chain (oracle/torch_ops.deform_mask follows generator_component4_15.py:148-173,
204-234, 294, 313-327 op for op).  fp32, tolerance 1e-5 relative (north_star) forward, 1e-4 for the gradients
(they are sums over N points accumulated in a different order)."""
import numpy as np
import pytest
import torch

from adaptpoint_b200 import imitator
from adaptpoint_b200.synth import make_clouds
from oracle import torch_ops

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("torch_backend_default")]


def _inputs(B, N, M, seed, with_mask):
    g = torch.Generator().manual_seed(seed)
    xyz = make_clouds(B, N, "shell", seed)
    anchor = xyz[:, torch.randperm(N, generator=g)[:M]].contiguous()
    degree = (torch.rand(B, M, 3, generator=g) - 0.5) * 0.35
    scale = torch.rand(B, M, 3, generator=g) * 2 + 1
    trl = (torch.rand(B, M, 3, generator=g) - 0.5) * 0.5
    affine = imitator._affine(degree, scale, trl)
    axis = ((torch.randint(1, 8, (B, 1), generator=g)[:, :, None] & (1 << torch.arange(3))) > 0).float().reshape(B, 3)
    mask = (torch.rand(B, N, generator=g) > 0.3).float() if with_mask else None
    return xyz, anchor, affine, axis, mask


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


@pytest.mark.parametrize("B,N,M,with_mask", [(2, 512, 4, True), (32, 2048, 4, True), (3, 1000, 1, False),
                                              (2, 4097, 8, True), (1, 20000, 4, False), (2, 31, 3, True)])
def test_deform_forward_backward(cuda_dev, B, N, M, with_mask):
    xyz, anchor, affine, axis, mask = _inputs(B, N, M, N + M, with_mask)
    w = torch.randn(B, N, 3, generator=torch.Generator().manual_seed(1))
    a_ref = affine.clone().requires_grad_(True)
    m_ref = mask.clone().requires_grad_(True) if with_mask else None
    want = torch_ops.deform_mask(xyz, anchor, a_ref, axis, m_ref, 0.5)
    (want * w).sum().backward()

    d = cuda_dev
    a_gpu = affine.to(d).requires_grad_(True)
    m_gpu = mask.to(d).requires_grad_(True) if with_mask else None
    got = imitator.deform_mask(xyz.to(d), anchor.to(d), a_gpu, axis.to(d), m_gpu, 0.5)
    (got * w.to(d)).sum().backward()
    assert _rel(got.detach().cpu().numpy(), want.detach().numpy()) < 1e-5
    assert _rel(a_gpu.grad.cpu().numpy(), a_ref.grad.numpy()) < 1e-4
    if with_mask:
        assert _rel(m_gpu.grad.cpu().numpy(), m_ref.grad.numpy()) < 1e-4
