"""TEST INFRASTRUCTURE ONLY -- CPU stand-ins for adaptpoint_b200.ops, used by
  * tests/ (host logic of the modules on a box without a GPU, and CPU-vs-GPU logits / gradient parity),
  * bench.py's `cpu_baseline` leg and `--impl reference` arm (BASELINE.json: "the reference's forward/backward
    run through a pure-PyTorch CPU equivalent of the same ops on the GPU box's own host cores").
Nothing under adaptpoint_b200/ imports this file.

Index-producing ops (FPS, ball query, 3-NN, kNN) come from liboracle.so (oracle/pointnet2_oracle.c: the plain-C
restatement of the reference kernels, bit-exact incl. tie rules), one cloud per worker thread (ctypes drops the
GIL); `pure_torch=True` switches them to torch-only formulations (argmax / masked sort / topk) whose indices
agree away from exact ties.  Differentiable ops (grouping, gather, interpolate) are torch.gather formulations
whose backward is autograd's, i.e. the same math as the reference's atomicAdd scatters
(group_points_gpu.cu:14-31, sampling_gpu.cu:53-70, interpolate_gpu.cu:127-149).
"""
import contextlib
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

from . import cpu as _orc

_POOL = None
PURE_TORCH = False


def _pool():
    global _POOL
    if _POOL is None:
        _POOL = ThreadPoolExecutor(max_workers=os.cpu_count() or 1)
    return _POOL


def _per_cloud(fn, B):
    if B == 1:
        return [fn(0)]
    return list(_pool().map(fn, range(B)))


@torch.no_grad()
def furthest_point_sample(xyz, npoint):
    """subsample.py:76-105 semantics on CPU."""
    x = xyz.detach().contiguous().float()
    B, N, _ = x.shape
    if PURE_TORCH:
        idx = torch.zeros(B, npoint, dtype=torch.long)
        dist = torch.full((B, N), 1e10)
        far = torch.zeros(B, dtype=torch.long)
        ar = torch.arange(B)
        for j in range(1, npoint):
            c = x[ar, far].unsqueeze(1)
            dist = torch.minimum(dist, ((x - c) ** 2).sum(-1))
            far = dist.argmax(-1)
            idx[:, j] = far
        return idx.int()
    xn = x.numpy()
    out = _per_cloud(lambda b: _orc.fps(xn[b:b + 1], npoint), B)
    return torch.from_numpy(np.concatenate(out, 0))


@torch.no_grad()
def ball_query(radius, nsample, xyz, new_xyz):
    """group.py:177-203 semantics on CPU: first `nsample` in-radius indices, padded with the first."""
    x, q = xyz.detach().contiguous().float(), new_xyz.detach().contiguous().float()
    B, N, _ = x.shape
    if PURE_TORCH:
        d2 = ((q.unsqueeze(2) - x.unsqueeze(1)) ** 2).sum(-1)                     # (B,M,N)
        key = torch.where(d2 < radius * radius, torch.arange(N).expand_as(d2), torch.full_like(d2, N, dtype=torch.long))
        first = key.sort(dim=-1).values[..., :nsample]
        pad = first[..., :1].clone()
        pad[pad == N] = 0
        return torch.where(first == N, pad.expand_as(first), first).int()
    xn, qn = x.numpy(), q.numpy()
    out = _per_cloud(lambda b: _orc.ball_query(radius, nsample, xn[b:b + 1], qn[b:b + 1]), B)
    return torch.from_numpy(np.concatenate(out, 0))


@torch.no_grad()
def knn_query(k, support, query, return_dist2=False):
    s, q = support.detach().contiguous().float(), query.detach().contiguous().float()
    B = s.shape[0]
    sn, qn = s.numpy(), q.numpy()
    out = _per_cloud(lambda b: _orc.knn(k, sn[b:b + 1], qn[b:b + 1]), B)
    idx = torch.from_numpy(np.concatenate([o[0] for o in out], 0))
    if return_dist2:
        return idx, torch.from_numpy(np.concatenate([o[1] for o in out], 0))
    return idx


def grouping_operation(features, idx):
    """group.py:120-137 (the reference's own torch_grouping_operation)."""
    B, C, _ = features.shape
    _, M, K = idx.shape
    flat = idx.reshape(B, 1, M * K).long().expand(-1, C, -1)
    return features.gather(2, flat).reshape(B, C, M, K)


torch_grouping_operation = grouping_operation


def group_minmax(features, idx):
    """CPU stand-in of ops.group_minmax: the unfused composition."""
    g = grouping_operation(features, idx)
    return g.max(dim=3)[0], g.min(dim=3)[0]


def gather_operation(features, idx):
    return features.gather(2, idx.long().unsqueeze(1).expand(-1, features.shape[1], -1))


@torch.no_grad()
def three_nn(unknown, known):
    u, k = unknown.detach().contiguous().float(), known.detach().contiguous().float()
    B = u.shape[0]
    if PURE_TORCH:
        d2 = ((u.unsqueeze(2) - k.unsqueeze(1)) ** 2).sum(-1)
        v, i = d2.topk(3, dim=-1, largest=False)
        return v.sqrt(), i.int()
    un, kn = u.numpy(), k.numpy()
    out = _per_cloud(lambda b: _orc.three_nn(un[b:b + 1], kn[b:b + 1]), B)
    d2 = torch.from_numpy(np.concatenate([o[0] for o in out], 0))
    idx = torch.from_numpy(np.concatenate([o[1] for o in out], 0))
    return d2.sqrt(), idx


def three_interpolate(features, idx, weight):
    B, C, m = features.shape
    n = idx.shape[1]
    g = features.gather(2, idx.reshape(B, 1, n * 3).long().expand(-1, C, -1)).reshape(B, C, n, 3)
    return (g * weight.unsqueeze(1)).sum(-1)


def three_interpolation(unknown_xyz, known_xyz, know_feat):
    dist, idx = three_nn(unknown_xyz, known_xyz)
    dist_recip = 1.0 / (dist + 1e-8)
    weight = dist_recip / dist_recip.sum(dim=2, keepdim=True)
    return three_interpolate(know_feat, idx, weight)


def fps(data, number):
    fps_idx = furthest_point_sample(data[:, :, :3].contiguous(), number)
    return torch.gather(data, 1, fps_idx.unsqueeze(-1).long().expand(-1, -1, data.shape[-1]))


def deform_mask(xyz, anchor, affine, proj_axis, mask=None, sigma=0.5):
    """Torch restatement of the deformation the fused kernel replaces, op for op as the reference writes it:
    local transform applied around each anchor (generator_component4_15.py:294, 167), kernel_regression
    (:204-234), normalize (:313-327), mask multiply (:173).  Differentiable by autograd."""
    B, N, _ = xyz.shape
    M = anchor.shape[1]
    A = affine[..., :9].reshape(B, M, 3, 3)
    trl = affine[..., 9:]
    pos_normalize = xyz.unsqueeze(1).repeat(1, M, 1, 1) - anchor.unsqueeze(-2)          # (B,M,N,3)
    transformed = pos_normalize @ A + trl.reshape(B, M, 1, 3) + anchor.reshape(B, M, 1, 3)
    sub = anchor.unsqueeze(-2).repeat(1, 1, N, 1) - xyz.unsqueeze(1).repeat(1, M, 1, 1)
    projection = proj_axis.reshape(B, 1, 3).float().unsqueeze(-2) * torch.eye(3)
    sub = torch.sqrt(((sub @ projection) ** 2).sum(dim=-1))
    weight = torch.exp(-0.5 * (sub ** 2) / (sigma ** 2))
    pos_new = (weight.unsqueeze(-1).repeat(1, 1, 1, 3) * transformed).sum(dim=1)
    pos_new = pos_new / weight.sum(dim=1, keepdims=True).transpose(1, 2).contiguous()
    pos = pos_new - pos_new.mean(axis=-2, keepdims=True)
    scale = (1 / torch.sqrt((pos ** 2).sum(dim=-1)).max(dim=-1)[0]) * 0.999999
    pos = scale.reshape(B, 1, 1).repeat(1, N, 3) * pos
    return pos if mask is None else pos * mask.unsqueeze(-1)


_NAMES = ["furthest_point_sample", "ball_query", "knn_query", "grouping_operation", "torch_grouping_operation",
          "group_minmax",
          "gather_operation", "three_nn", "three_interpolate", "three_interpolation", "fps"]


@contextlib.contextmanager
def patched(pure_torch=False):
    """Temporarily point adaptpoint_b200.ops at the CPU stand-ins (tests / CPU baseline only).  Modules must be
    constructed AND run inside the context (SetAbstraction binds its sampler at construction)."""
    global PURE_TORCH
    import sys
    from adaptpoint_b200 import ops
    me = sys.modules[__name__]
    from adaptpoint_b200 import imitator
    saved = {n: getattr(ops, n) for n in _NAMES}
    saved_deform = imitator.deform_mask
    saved_flag = PURE_TORCH
    PURE_TORCH = pure_torch
    try:
        for n in _NAMES:
            setattr(ops, n, getattr(me, n))
        imitator.deform_mask = deform_mask
        yield
    finally:
        for n, v in saved.items():
            setattr(ops, n, v)
        imitator.deform_mask = saved_deform
        PURE_TORCH = saved_flag
