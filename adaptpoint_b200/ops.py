"""The reference's pointnet2_batch operator API over libapb200.so.

Same names, argument order, dtypes and layouts as
  openpoints/models/layers/subsample.py:76-156   (FurthestPointSampling, GatherOperation, fps)
  openpoints/models/layers/group.py:76-203       (GroupingOperation, GatherOperation, BallQuery)
  openpoints/models/layers/upsampling.py:11-102  (ThreeNN, ThreeInterpolate, three_interpolation)
Differences: outputs are allocated on the INPUT's device (the reference's legacy
`torch.cuda.FloatTensor(...)` ctors use the current device and ignore `device=`), kernels run on
torch's current stream, and errors raise instead of exit(-1).  There is no CPU path: CPU tensors
raise `ApbError`.
"""
from typing import Tuple

import torch
from torch.autograd import Function

from . import _cabi


def _check(t, name, dtype=None):
    if not t.is_cuda:
        raise _cabi.ApbError(f"{name}: CUDA tensor required (adaptpoint_b200 has no CPU path)")
    assert t.is_contiguous(), f"{name} must be contiguous"
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"{name}: expected {dtype}, got {t.dtype}")


class FurthestPointSampling(Function):
    @staticmethod
    def forward(ctx, xyz: torch.Tensor, npoint: int) -> torch.Tensor:
        """xyz (B,N,3) f32 -> idx (B,npoint) int32; reference subsample.py:77-100."""
        _check(xyz, "xyz", torch.float32)
        B, N, _ = xyz.size()
        output = torch.empty((B, npoint), dtype=torch.int32, device=xyz.device)
        # temp=None: the kernel keeps the running min-distance (the reference's 1e10-filled scratch,
        # subsample.py:94) in registers and never spills it to HBM.
        _cabi.call("apb_fps_f32", [xyz, None, output, B, N, int(npoint)], xyz.device)
        ctx.mark_non_differentiable(output)
        return output

    @staticmethod
    def backward(ctx, a=None):
        return None, None


furthest_point_sample = FurthestPointSampling.apply


class GatherOperation(Function):
    @staticmethod
    def forward(ctx, features: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        """features (B,C,N), idx (B,npoint) int32 -> (B,C,npoint); reference subsample.py:108-129."""
        _check(features, "features", torch.float32)
        _check(idx, "idx", torch.int32)
        B, npoint = idx.size()
        _, C, N = features.size()
        output = torch.empty((B, C, npoint), dtype=torch.float32, device=features.device)
        _cabi.call("apb_gather_f32", [features, idx, output, B, C, N, npoint], features.device)
        ctx.for_backwards = (idx, C, N)
        return output

    @staticmethod
    def backward(ctx, grad_out):
        idx, C, N = ctx.for_backwards
        B, npoint = idx.size()
        grad_features = torch.zeros((B, C, N), dtype=torch.float32, device=grad_out.device)
        grad_out_data = grad_out.detach().contiguous().float()
        _cabi.call("apb_gather_grad_f32", [grad_out_data, idx, grad_features, B, C, N, npoint], grad_out.device)
        return grad_features, None


gather_operation = GatherOperation.apply


# Support clouds of at least this many points go through the cell grid (below it the tiled O(M*N) kernel is as fast: its
# whole cloud sits in one or two shared-memory tiles and the grid costs four small build launches).
BALL_QUERY_GRID_MIN_N = 4096


class BallQuery(Function):
    @staticmethod
    def forward(ctx, radius: float, nsample: int, xyz: torch.Tensor, new_xyz: torch.Tensor) -> torch.Tensor:
        """xyz (B,N,3) support, new_xyz (B,npoint,3) centres -> idx (B,npoint,nsample) int32;
        reference group.py:177-196 (note its wrapper takes new_xyz before xyz)."""
        _check(new_xyz, "new_xyz", torch.float32)
        _check(xyz, "xyz", torch.float32)
        B, N, _ = xyz.size()
        npoint = new_xyz.size(1)
        idx = torch.empty((B, npoint, nsample), dtype=torch.int32, device=xyz.device)
        if N >= BALL_QUERY_GRID_MIN_N and 0 < nsample <= 96 and B <= 65535:
            # cell-grid kernel (csrc/query_grid.cu): same rows bit for bit, O(M * neighbourhood) pair evaluations
            nbytes = _cabi.load().apb_ball_query_grid_ws_bytes(B, N)
            ws = torch.empty(nbytes, dtype=torch.uint8, device=xyz.device)
            _cabi.call("apb_ball_query_grid_f32", [new_xyz, xyz, idx, ws, nbytes, B, N, npoint, float(radius),
                                                   int(nsample)], xyz.device)
        else:
            _cabi.call("apb_ball_query_f32", [new_xyz, xyz, idx, B, N, npoint, float(radius), int(nsample)], xyz.device)
        ctx.mark_non_differentiable(idx)
        return idx

    @staticmethod
    def backward(ctx, a=None):
        return None, None, None, None


ball_query = BallQuery.apply


class GroupingOperation(Function):
    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, features: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        """features (B,C,N), idx (B,npoint,nsample) int32 -> (B,C,npoint,nsample); group.py:76-98."""
        _check(features, "features", torch.float32)
        _check(idx, "idx", torch.int32)
        B, nfeatures, nsample = idx.size()
        _, C, N = features.size()
        output = torch.empty((B, C, nfeatures, nsample), dtype=torch.float32, device=features.device)
        _cabi.call("apb_group_f32", [features, idx, output, B, C, N, nfeatures, nsample], features.device)
        ctx.for_backwards = (idx, N)
        return output

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, grad_out: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        idx, N = ctx.for_backwards
        B, C, npoint, nsample = grad_out.size()
        grad_features = torch.zeros((B, C, N), dtype=torch.float32, device=grad_out.device)
        grad_out_data = grad_out.detach().contiguous().float()
        _cabi.call("apb_group_grad_f32", [grad_out_data, idx, grad_features, B, C, N, npoint, nsample],
                   grad_out.device)
        return grad_features, None


grouping_operation = GroupingOperation.apply


class GroupMinMax(Function):
    @staticmethod
    def forward(ctx, features: torch.Tensor, idx: torch.Tensor):
        """features (B,C,N), idx (B,npoint,nsample) int32 -> (max, min) over the nsample neighbours, each (B,C,npoint):
        grouping_operation(features, idx).max(-1)[0] / .min(-1)[0] without the (B,C,npoint,nsample) tensor
        (apb_group_minmax_f32).  The backward pass scatters (B,C,npoint) gradients to the arg-max / arg-min points."""
        _check(features, "features", torch.float32)
        _check(idx, "idx", torch.int32)
        B, M, K = idx.size()
        _, C, N = features.size()
        dev = features.device
        vmax = torch.empty((B, C, M), dtype=torch.float32, device=dev)
        vmin = torch.empty((B, C, M), dtype=torch.float32, device=dev)
        amax = torch.empty((B, C, M), dtype=torch.int32, device=dev)
        amin = torch.empty((B, C, M), dtype=torch.int32, device=dev)
        _cabi.call("apb_group_minmax_f32", [features, idx, vmax, vmin, amax, amin, B, C, N, M, K], dev)
        ctx.save_for_backward(amax, amin)
        ctx.N = N
        return vmax, vmin

    @staticmethod
    def backward(ctx, gmax, gmin):
        amax, amin = ctx.saved_tensors
        B, C, M = amax.shape
        grad = torch.zeros((B, C, ctx.N), dtype=torch.float32, device=amax.device)
        if gmax is not None:
            grad.scatter_add_(2, amax.long(), gmax.detach().float().contiguous())
        if gmin is not None:
            grad.scatter_add_(2, amin.long(), gmin.detach().float().contiguous())
        return grad, None


def group_minmax(features, idx):
    """(max, min) over the neighbours of grouping_operation(features, idx); falls back to the unfused composition of
    our own kernels when a feature row does not fit in shared memory (N > 51200)."""
    if features.shape[2] * 4 > 200 * 1024:
        g = grouping_operation(features, idx)
        return g.max(dim=3)[0], g.min(dim=3)[0]
    return GroupMinMax.apply(features, idx)


class SelfAttention16(Function):
    @staticmethod
    def forward(ctx, q, k, v, heads: int):
        """q,k,v (B,M,heads*16) f32 token-major -> softmax(q k^T / 4) v, (B,M,heads*16); scores stay in registers
        (apb_attn_fwd_f32; reference generator_component4_15.py:434-480)."""
        q, k, v = (t.detach().contiguous().float() for t in (q, k, v))
        B, M, C = q.shape
        o = torch.empty_like(q)
        lse = torch.empty((B, heads, M), dtype=torch.float32, device=q.device)
        _cabi.call("apb_attn_fwd_f32", [q, k, v, o, lse, B, M, heads, C // heads], q.device)
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.heads = heads
        return o

    @staticmethod
    def backward(ctx, grad_out):
        q, k, v, o, lse = ctx.saved_tensors
        B, M, C = q.shape
        go = grad_out.detach().contiguous().float()
        dq, dk, dv = torch.empty_like(q), torch.empty_like(q), torch.empty_like(q)
        delta = torch.empty_like(lse)
        _cabi.call("apb_attn_bwd_f32", [q, k, v, o, go, lse, dq, dk, dv, delta, B, M, ctx.heads, C // ctx.heads],
                   q.device)
        return dq, dk, dv, None


class SelfAttentionSmall(Function):
    @staticmethod
    def forward(ctx, q, k, v, heads: int):
        """Few tokens (M <= 32), any head width (apb_attn_small_fwd_f32): the imitator's anchor head."""
        q, k, v = (t.detach().contiguous().float() for t in (q, k, v))
        B, M, C = q.shape
        o = torch.empty_like(q)
        p = torch.empty((B, heads, M, M), dtype=torch.float32, device=q.device)
        _cabi.call("apb_attn_small_fwd_f32", [q, k, v, o, p, B, M, heads, C // heads], q.device)
        ctx.save_for_backward(q, k, v, p)
        ctx.heads = heads
        return o

    @staticmethod
    def backward(ctx, grad_out):
        q, k, v, p = ctx.saved_tensors
        B, M, C = q.shape
        go = grad_out.detach().contiguous().float()
        dq, dk, dv = torch.empty_like(q), torch.empty_like(q), torch.empty_like(q)
        _cabi.call("apb_attn_small_bwd_f32", [q, k, v, p, go, dq, dk, dv, B, M, ctx.heads, C // ctx.heads], q.device)
        return dq, dk, dv, None


def self_attention(q, k, v, heads):
    """Multi-head self-attention softmax(q k^T / sqrt(d)) v over (B,M,C) token-major tensors: the flash-style kernel for
    head_dim 16 (any M), the few-token kernel for M <= 32 (any head_dim).  Nothing else occurs on the path, and there
    is no library fallback."""
    if q.shape[-1] // heads == 16:
        return SelfAttention16.apply(q, k, v, heads)
    if q.shape[1] <= 32:
        return SelfAttentionSmall.apply(q, k, v, heads)
    raise _cabi.ApbError(f"self_attention: head_dim {q.shape[-1] // heads} with {q.shape[1]} tokens is outside the "
                         "kernel family (head_dim 16, or <= 32 tokens)")


def grouping_operation_bf16(features, idx):
    """grouping_operation for bf16 features (no autograd: the bf16 shared-MLP path is inference-only):
    features (B,C,N) bf16, idx (B,M,K) int32 -> (B,C,M,K) bf16, a bit-exact copy through apb_group_bf16."""
    _check(features, "features", torch.bfloat16)
    _check(idx, "idx", torch.int32)
    B, C, N = features.shape
    _, M, K = idx.shape
    out = torch.empty((B, C, M, K), dtype=torch.bfloat16, device=features.device)
    _cabi.call("apb_group_bf16", [features, idx, out, B, C, N, M, K], features.device)
    return out


def torch_grouping_operation(features, idx):
    """Pure-torch grouping kept for API parity (reference group.py:120-137)."""
    all_idx = idx.reshape(idx.shape[0], -1)
    all_idx = all_idx.unsqueeze(1).repeat(1, features.shape[1], 1)
    grouped_features = features.gather(2, all_idx.long())
    return grouped_features.reshape(idx.shape[0], features.shape[1], idx.shape[1], idx.shape[2])


class ThreeNN(Function):
    @staticmethod
    def forward(ctx, unknown: torch.Tensor, known: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """unknown (B,n,3), known (B,m,3) -> (dist (B,n,3) [sqrt applied], idx (B,n,3) int32);
        reference upsampling.py:11-33."""
        _check(unknown, "unknown", torch.float32)
        _check(known, "known", torch.float32)
        B, N, _ = unknown.size()
        m = known.size(1)
        dist2 = torch.empty((B, N, 3), dtype=torch.float32, device=unknown.device)
        idx = torch.empty((B, N, 3), dtype=torch.int32, device=unknown.device)
        _cabi.call("apb_three_nn_f32", [unknown, known, dist2, idx, B, N, m], unknown.device)
        ctx.mark_non_differentiable(idx)
        return torch.sqrt(dist2), idx

    @staticmethod
    def backward(ctx, a=None, b=None):
        return None, None


three_nn = ThreeNN.apply


class ThreeInterpolate(Function):
    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, features: torch.Tensor, idx: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
        """features (B,C,m), idx (B,n,3) int32, weight (B,n,3) -> (B,C,n); upsampling.py:43-67."""
        _check(features, "features", torch.float32)
        _check(idx, "idx", torch.int32)
        _check(weight, "weight", torch.float32)
        B, c, m = features.size()
        n = idx.size(1)
        ctx.three_interpolate_for_backward = (idx, weight, m)
        output = torch.empty((B, c, n), dtype=torch.float32, device=features.device)
        _cabi.call("apb_three_interp_f32", [features, idx, weight, output, B, c, m, n], features.device)
        return output

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, grad_out: torch.Tensor):
        idx, weight, m = ctx.three_interpolate_for_backward
        B, c, n = grad_out.size()
        grad_features = torch.zeros((B, c, m), dtype=torch.float32, device=grad_out.device)
        grad_out_data = grad_out.detach().contiguous().float()
        _cabi.call("apb_three_interp_grad_f32", [grad_out_data, idx, weight, grad_features, B, c, n, m],
                   grad_out.device)
        return grad_features, None, None


three_interpolate = ThreeInterpolate.apply


def three_interpolation(unknown_xyz, known_xyz, know_feat):
    """reference upsampling.py:92-102: inverse-distance weights over the 3 nearest known points."""
    dist, idx = three_nn(unknown_xyz, known_xyz)
    dist_recip = 1.0 / (dist + 1e-8)
    norm = torch.sum(dist_recip, dim=2, keepdim=True)
    weight = dist_recip / norm
    return three_interpolate(know_feat, idx, weight)


def fps(data, number):
    """data (B,N,C>=3) -> (B,number,C) sampled rows; reference subsample.py:148-156."""
    fps_idx = furthest_point_sample(data[:, :, :3].contiguous(), number)
    return torch.gather(data, 1, fps_idx.unsqueeze(-1).long().expand(-1, -1, data.shape[-1]))


@torch.no_grad()
def knn_query(k: int, support: torch.Tensor, query: torch.Tensor, return_dist2: bool = False):
    """k nearest support points of every query, ordered by (squared distance, index).
    support (B,N,3), query (B,M,3) -> idx (B,M,k) int32 [, dist2 (B,M,k)].
    Stands in for torch.cdist + topk in the reference's KNN module (group.py:12-28)."""
    _check(support, "support", torch.float32)
    _check(query, "query", torch.float32)
    B, N, _ = support.size()
    M = query.size(1)
    idx = torch.empty((B, M, k), dtype=torch.int32, device=support.device)
    d2 = torch.empty((B, M, k), dtype=torch.float32, device=support.device) if return_dist2 else None
    _cabi.call("apb_knn_f32", [query, support, idx, d2, B, N, M, int(k)], support.device)
    return (idx, d2) if return_dist2 else idx
