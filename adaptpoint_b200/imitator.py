"""The AdaptPoint imitator (generator), its discriminator and PointWOLF, over adaptpoint_b200.ops.

Reference:
  AdaptPoint_Augmentor / SAComponent / PointsetGrouper / PointNetFeaturePropagation / Producefactor /
  Anchor_selfattention      openpoints/models_adaptpoint/generator_component4_15.py:119-718
  PointDiscriminator1       openpoints/models_adaptpoint/point_discriminator.py:17-73, 149-191
  PointWOLF_classversion    openpoints/online_aug/pointwolf.py:14-178
Module / parameter names follow the reference so `model_gan.pth` checkpoints (train_autoaug.py:71-78) load.

What runs where:
  * FPS, ball query, grouping (fwd+bwd), kNN, 3-NN, 3-interpolation  -> libapb200.so kernels (ops.py);
  * the per-point deformation (local affine transforms, kernel regression, re-normalisation, mask)
    -> ONE fused kernel, apb_deform_mask_{fwd,bwd}_f32 (csrc/deform.cu), via `deform_mask` below;
  * 1x1 convs / BatchNorm / attention matmuls -> torch (library GEMMs; SURVEY.md 8f ranks them "next").
Random draws keep the reference's generators and order (dropout / axes from the CPU generator,
generator_component4_15.py:245-247,308; Gumbel noise on the device, :714) and can be injected through
`noise=` so tests are reproducible across machines.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.autograd import Function

from . import _cabi, geometry, ops
from .norm import ConvBlock, TcBatchNorm1d
from .spectral import spectral_norm
from .tcconv import TcConv1d, TcConv2d, conv_max


# ----------------------------------------------------------------------------------------------------------
# fused deformation op
# ----------------------------------------------------------------------------------------------------------
class DeformMask(Function):
    """out = mask * normalise(kernel_regression(local affine transforms)); grads for `affine` and `mask`."""

    @staticmethod
    def forward(ctx, xyz, anchor, affine, proj_axis, mask, sigma):
        for t, name in ((xyz, "xyz"), (anchor, "anchor"), (affine, "affine"), (proj_axis, "proj_axis")):
            ops._check(t, name, torch.float32)
        B, N, _ = xyz.shape
        M = anchor.shape[1]
        if mask is not None:
            ops._check(mask, "mask", torch.float32)
        out = torch.empty_like(xyz)
        stats = torch.empty((B, 8), dtype=torch.float32, device=xyz.device)
        _cabi.call("apb_deform_mask_fwd_f32", [xyz, anchor, affine, proj_axis, mask, float(sigma), out, stats,
                                               B, N, M], xyz.device)
        ctx.save_for_backward(xyz, anchor, affine, proj_axis, mask, stats)
        ctx.sigma = float(sigma)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        xyz, anchor, affine, proj_axis, mask, stats = ctx.saved_tensors
        B, N, _ = xyz.shape
        M = anchor.shape[1]
        g = grad_out.detach().contiguous().float()
        grad_affine = torch.empty_like(affine)
        grad_mask = torch.empty((B, N), dtype=torch.float32, device=xyz.device) if mask is not None else None
        _cabi.call("apb_deform_mask_bwd_f32", [xyz, anchor, affine, proj_axis, mask, ctx.sigma, stats, g,
                                               grad_affine, grad_mask, B, N, M], xyz.device)
        return None, None, grad_affine, None, grad_mask, None


def deform_mask(xyz, anchor, affine, proj_axis, mask=None, sigma=0.5):
    """xyz (B,N,3), anchor (B,M,3), affine (B,M,12) [A = R S row-major, then trl], proj_axis (B,3),
    mask (B,N) or None -> (B,N,3)."""
    return DeformMask.apply(xyz.contiguous(), anchor.contiguous(), affine.contiguous(),
                            proj_axis.contiguous().float(), None if mask is None else mask.contiguous(), sigma)


def _random_axis(batch, n_axis):
    """(batch,n_axis,3) 0/1 axes from the CPU generator: bit j of randint(1,8) (generator_component4_15.py:299-311)."""
    axis = torch.randint(1, 8, (batch, n_axis))
    return ((axis[:, :, None] & (1 << torch.arange(3))) > 0).int()


def _rotation(degree):
    """The reference's 9-term matrix, including its [1][1] = sz*sy*sx + cz*cy term
    (generator_component4_15.py:285-293, pointwolf.py:137-145); degree (B,M,3) -> (B,M,3,3)."""
    sin, cos = torch.sin(degree), torch.cos(degree)
    sx, sy, sz = sin[..., 0], sin[..., 1], sin[..., 2]
    cx, cy, cz = cos[..., 0], cos[..., 1], cos[..., 2]
    rows = [cz * cy, cz * sy * sx - sz * cx, cz * sy * cx + sz * sx,
            sz * cy, sz * sy * sx + cz * cy, sz * sy * cx - cz * sx,
            -sy, cy * sx, cy * cx]
    return torch.stack(rows, dim=-1).reshape(*degree.shape[:-1], 3, 3)


def _affine(degree, scale, trl):
    """(B,M,12): A = R @ diag(scale) row-major, then trl."""
    A = _rotation(degree) * scale.unsqueeze(-2)
    return torch.cat([A.reshape(*A.shape[:-2], 9), trl], dim=-1).contiguous()


def _index_points(points, idx):
    """points (B,N,C), idx (B,S) -> (B,S,C)."""
    return torch.gather(points, 1, idx.unsqueeze(-1).expand(-1, -1, points.shape[-1]))


# ----------------------------------------------------------------------------------------------------------
# PointWOLF (random parameters, no mask)
# ----------------------------------------------------------------------------------------------------------
class PointWOLF_classversion(object):
    def __init__(self, w_num_anchor=4, w_sigma=0.5, w_R_range=10, w_S_range=3, w_T_range=0.25):
        self.num_anchor, self.sigma = w_num_anchor, w_sigma
        self.R_range = (-abs(w_R_range), abs(w_R_range))
        self.S_range = (1., w_S_range)
        self.T_range = (-abs(w_T_range), abs(w_T_range))

    def draw(self, B):
        """CPU-generator draws in the reference's order (pointwolf.py:113-131, 96)."""
        M = self.num_anchor
        dropout = torch.bernoulli(torch.Tensor(B, M, 3).uniform_(0, 1))
        axis = _random_axis(B, M)
        r = torch.FloatTensor(B, M, 3).uniform_(*self.R_range)
        s = torch.FloatTensor(B, M, 3).uniform_(*self.S_range)
        t = torch.FloatTensor(B, M, 3).uniform_(*self.T_range)
        proj_axis = _random_axis(B, 1)
        return dict(dropout=dropout, axis=axis, R=r, S=s, T=t, proj_axis=proj_axis)

    @torch.no_grad()
    def __call__(self, xyz, noise=None):
        B, N, _ = xyz.shape
        dev = xyz.device
        noise = self.draw(B) if noise is None else noise
        fps_idx, anchor = geometry.sample(xyz, self.num_anchor)
        drop, axis = noise["dropout"].to(dev), noise["axis"].to(dev).float()
        degree = math.pi * noise["R"].to(dev) / 180.0 * drop[:, :, 0:1]
        scale = noise["S"].to(dev) * drop[:, :, 1:2] * axis
        scale = scale + 1 * (scale == 0)
        trl = noise["T"].to(dev) * drop[:, :, 2:3] * axis
        out = deform_mask(xyz, anchor, _affine(degree, scale, trl), noise["proj_axis"].to(dev).float().reshape(B, 3),
                          None, self.sigma)
        return xyz, out


# ----------------------------------------------------------------------------------------------------------
# imitator network
# ----------------------------------------------------------------------------------------------------------
class ConvBNReLU1D(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size=1, bias=True, activation="relu"):
        super().__init__()
        if activation.lower() != "relu":
            raise NotImplementedError(activation)
        self.act = nn.ReLU(inplace=True)
        self.net = ConvBlock(TcConv1d(in_channels, out_channels, kernel_size, bias=bias),
                             TcBatchNorm1d(out_channels), self.act)

    def forward(self, x):
        return self.net(x)


class PointNetFeaturePropagation(nn.Module):
    """3-NN inverse-distance interpolation + skip concat + Conv/BN/ReLU (generator_component4_15.py:330-365)."""

    def __init__(self, in_channel, out_channel, blocks=1, groups=1, res_expansion=1.0, bias=False, activation="relu"):
        super().__init__()
        self.fuse = ConvBNReLU1D(in_channel, out_channel, 1, bias=bias)

    def forward(self, xyz1, xyz2, points1, points2):
        up = ops.three_interpolation(xyz1.contiguous(), xyz2.contiguous(), points2.contiguous())
        return self.fuse(up if points1 is None else torch.cat([points1, up], dim=1))


class PointsetGrouper(nn.Module):
    """FPS -> ball query -> group -> (x - anchor feature) * alpha + beta -> max over neighbours
    (generator_component4_15.py:368-431).  Features stay channel-major (B,C,N) so grouping is one
    apb_group_f32 launch instead of the reference's advanced-indexing gather of a (B,np,k,C) tensor."""

    def __init__(self, channel, reduce, kneighbors, radi, normalize="anchor", **kwargs):
        super().__init__()
        self.reduce, self.kneighbors, self.radi = reduce, kneighbors, radi
        self.normalize = normalize.lower() if normalize is not None else None
        if self.normalize not in ("center", "anchor"):
            self.normalize = None
        if self.normalize is not None:
            self.affine_alpha = nn.Parameter(torch.ones([1, 1, 1, channel]))
            self.affine_beta = nn.Parameter(torch.zeros([1, 1, 1, channel]))

    def forward(self, xyz, feats):
        """xyz (B,N,3), feats (B,C,N) -> new_xyz (B,np,3), new_feats (B,C,np)."""
        xyz = xyz.contiguous()
        B, C, N = feats.shape
        fps_idx, new_xyz = geometry.sample(xyz, N // self.reduce)          # shared FPS chain when xyz belongs to a plan
        centre = torch.gather(feats, 2, fps_idx.unsqueeze(1).expand(-1, C, -1))            # (B,C,np)
        idx = geometry.ball_query(self.radi, self.kneighbors, xyz, new_xyz)
        if self.normalize == "center":   # needs the mean over the neighbours: unfused composition
            grouped = ops.grouping_operation(feats.contiguous(), idx)                        # (B,C,np,k)
            grouped = self.affine_alpha.view(1, C, 1, 1) * (grouped - grouped.mean(dim=3, keepdim=True)) \
                + self.affine_beta.view(1, C, 1, 1)
            return new_xyz, grouped.max(dim=3)[0]
        # alpha*(g - centre) + beta is monotone in g: its max over the neighbours is the same expression of the
        # max (alpha >= 0) or the min (alpha < 0) of g, bit for bit, so the (B,C,np,k) tensor is never built
        gmax, gmin = ops.group_minmax(feats.contiguous(), idx)                               # (B,C,np) each
        if self.normalize is None:
            return new_xyz, gmax
        alpha, beta = self.affine_alpha.view(1, C, 1), self.affine_beta.view(1, C, 1)
        sel = torch.where(alpha >= 0, gmax, gmin)
        return new_xyz, alpha * (sel - centre) + beta


class Anchor_selfattention(nn.Module):
    """Multi-head self-attention with an xyz positional embedding (generator_component4_15.py:434-480).
    softmax(q k^T / sqrt(d)) v runs on csrc/attention.cu (ops.self_attention), so the (B,heads,M,M) matrix of the mask
    branch (M = N = 2048) is never materialised."""

    def __init__(self, dim, head_num):
        super().__init__()
        self.dim, self.head_num = dim, head_num
        self.head_dim = int(dim // head_num)
        self.to_qkv = nn.Linear(dim, dim * 3, bias=False)
        self.pos_embedding = nn.Sequential(TcConv1d(3, dim, 1), TcBatchNorm1d(dim))
        self.res = nn.Sequential(TcConv1d(dim, dim, 1), TcBatchNorm1d(dim))

    def forward(self, x, xyz=None):
        """x (B,M,C), xyz (B,M,3) -> (B,M,C)."""
        B, M, C = x.shape
        rel = xyz - xyz.mean(dim=1, keepdim=True)
        emb = self.pos_embedding(rel.permute(0, 2, 1)).permute(0, 2, 1)
        q, k, v = self.to_qkv(x).chunk(3, dim=-1)
        if x.is_cuda:      # our fp32 kernels (flash-style for head_dim 16, few-token otherwise), token-major in / out
            o = ops.self_attention(q + emb, k + emb, v + emb, self.head_num)
        else:              # CPU tensors only: the op-patched CPU tests and the CPU baseline of bench.py
            split = lambda t: (t + emb).reshape(B, M, self.head_num, self.head_dim).permute(0, 2, 1, 3)
            o = F.scaled_dot_product_attention(split(q), split(k), split(v))
            o = o.permute(0, 2, 1, 3).reshape(B, M, -1)
        return self.res(o.permute(0, 2, 1)).permute(0, 2, 1)


class Producefactor(nn.Module):
    """Per-anchor R/S/T logits (generator_component4_15.py:534-585)."""

    def __init__(self, kneighbors, out_channels):
        super().__init__()
        self.keighbors, self.out_channels = kneighbors, out_channels
        self.global_layer = nn.Sequential(TcConv1d(3, out_channels, 1, bias=False), TcBatchNorm1d(out_channels))
        self.prob_head = nn.Sequential(TcConv1d(out_channels * 2, 9, 1, bias=False), TcBatchNorm1d(9))
        self.anchor_selfattention = Anchor_selfattention(dim=out_channels, head_num=4)

    def forward(self, a_points, sa_x, sa_xyz, xyz_raw=None):
        """a_points (B,A,3), sa_x (B,C,np) channel-major, sa_xyz (B,np,3) -> (B,A,9)."""
        A = a_points.shape[1]
        idx = ops.knn_query(self.keighbors, sa_xyz.contiguous(), a_points.contiguous())       # (B,A,k)
        local = ops.group_minmax(sa_x.contiguous(), idx)[0].permute(0, 2, 1)                  # (B,A,C)
        local = local + self.anchor_selfattention(x=local, xyz=a_points)
        glob = self.global_layer(a_points.permute(0, 2, 1)).permute(0, 2, 1).max(dim=1, keepdim=True)[0]
        feat = torch.cat([local, glob.repeat(1, A, 1)], dim=-1)
        return self.prob_head(feat.permute(0, 2, 1)).permute(0, 2, 1)


def gumbel_softmax_hard(logits, tau, gumbels=None, hard=None):
    """F.gumbel_softmax(hard=True) with injectable noise (straight-through one-hot).  `hard` (same shape, one-hot)
    replaces the arg-max decision -- parity tests feed the reference run's decisions so that a near-tie logit cannot
    flip a point between runs; the gradient still flows through the soft-max of this run."""
    if gumbels is None:
        gumbels = -torch.empty_like(logits, memory_format=torch.legacy_contiguous_format).exponential_().log()
    y_soft = ((logits + gumbels) / tau).softmax(-1)
    if hard is None:
        index = y_soft.max(-1, keepdim=True)[1]
        y_hard = torch.zeros_like(logits, memory_format=torch.legacy_contiguous_format).scatter_(-1, index, 1.0)
    else:
        y_hard = hard.to(logits.dtype)
    return y_hard - y_soft.detach() + y_soft


class SAComponent(nn.Module):
    """The imitator network: 4 SA stages, 4 FP stages, anchor head and the mask branch
    (generator_component4_15.py:588-718)."""

    def __init__(self, in_channel=3, embed_dim=64, res_expansion=1.0, activation="relu", bias=False,
                 normalize="anchor", dim_expansion=(2, 2, 2, 2), radii=(0.1, 0.2, 0.4, 0.8),
                 k_neighbors=(24, 24, 24, 24), reducers=(2, 2, 2, 2), **kwargs):
        super().__init__()
        self.stages = len(dim_expansion)
        self.embedding = ConvBNReLU1D(in_channel, embed_dim, bias=bias, activation=activation)
        self.extract_feat_list = nn.ModuleList()
        self.pointset_grouper_list = nn.ModuleList()
        last, channels = embed_dim, [embed_dim]
        for i in range(self.stages):
            out = last * dim_expansion[i]
            self.extract_feat_list.append(ConvBNReLU1D(last, out, 1, bias=bias, activation=activation))
            self.pointset_grouper_list.append(PointsetGrouper(channel=out, reduce=reducers[i],
                                                              kneighbors=k_neighbors[i], radi=radii[i],
                                                              normalize=normalize))
            last = out
            channels.append(out)
        self.head = Producefactor(kneighbors=24, out_channels=last)
        self.decode_list = nn.ModuleList([
            PointNetFeaturePropagation(channels[-(i + 1)] + channels[-(i + 2)], channels[-(i + 2)], bias=bias,
                                       activation=activation) for i in range(self.stages)])
        self.localfeat_mask_selfattention = Anchor_selfattention(dim=embed_dim, head_num=4)
        self.extract_local_feat_masking = nn.Sequential(TcConv1d(embed_dim, 3, 1, bias=False), TcBatchNorm1d(3))
        self.extract_global_feat_masking = nn.Sequential(TcConv1d(last, 3, 1, bias=False), TcBatchNorm1d(3))
        self.fuse_masking = nn.Sequential(TcConv1d(6, 2, 1, bias=False), TcBatchNorm1d(2))

    def forward(self, x, a_index=None, gumbels=None, hard_mask=None):
        """x (B,N,3), a_index (B,M) -> prob (B,M,9), masking (B,N,2) hard one-hot (straight-through)."""
        _, N, _ = x.shape
        a_points = _index_points(x, a_index)
        xyz = x
        f = self.embedding(x.permute(0, 2, 1).contiguous())
        xyz_list, x_list = [xyz], [f]
        for i in range(self.stages):
            f = self.extract_feat_list[i](f)
            xyz, f = self.pointset_grouper_list[i](xyz, f)
            xyz_list.append(xyz)
            x_list.append(f)
        for i in range(self.stages):
            x_list[-(i + 2)] = self.decode_list[i](xyz1=xyz_list[-(i + 2)], xyz2=xyz_list[-(i + 1)],
                                                   points1=x_list[-(i + 2)], points2=x_list[-(i + 1)])
        prob = self.head(a_points=a_points, sa_x=f, sa_xyz=xyz)

        local = x_list[0].permute(0, 2, 1)
        local = self.localfeat_mask_selfattention(x=local, xyz=xyz_list[0]) + local
        m_local = self.extract_local_feat_masking(local.permute(0, 2, 1))                    # (B,3,N)
        m_global = self.extract_global_feat_masking(x_list[-1]).max(dim=2, keepdim=True)[0]    # (B,3,1)
        masking = self.fuse_masking(torch.cat([m_local, m_global.repeat(1, 1, N)], dim=1)).permute(0, 2, 1)
        return prob, gumbel_softmax_hard(masking, 0.1, gumbels, hard_mask)


class AdaptPoint_Augmentor(nn.Module):
    """The imitator: predicts per-anchor deformation parameters and a per-point mask, applies them with the
    fused deformation kernel (generator_component4_15.py:119-327)."""

    def __init__(self, w_num_anchor=4, w_sigma=0.5, w_R_range=10, w_S_range=3, w_T_range=0.25):
        super().__init__()
        self.num_anchor, self.sigma = w_num_anchor, w_sigma
        self.w_R_range, self.w_S_range, self.w_T_range = w_R_range, w_S_range, w_T_range
        self.predict_prob_layer = SAComponent()

    def draw(self, B):
        """CPU-generator draws in the reference's order (local_transformaton :245-247, kernel_regression :221)."""
        dropout = torch.bernoulli(torch.Tensor(B, self.num_anchor, 3).uniform_(0, 1))
        axis = _random_axis(B, self.num_anchor)
        proj_axis = _random_axis(B, 1)
        return dict(dropout=dropout, axis=axis, proj_axis=proj_axis)

    def forward(self, xyz, noise=None):
        """xyz (B,N,3) -> (xyz, xyz_new).  noise: optional dict(dropout, axis, proj_axis[, gumbel][, hard_mask])."""
        B, N, _ = xyz.shape
        dev = xyz.device
        fps_idx, anchor = geometry.sample(xyz, self.num_anchor)
        on_dev = lambda t: None if t is None else t.to(dev)
        probs, masking = self.predict_prob_layer(xyz, fps_idx, None if noise is None else on_dev(noise.get("gumbel")),
                                                 None if noise is None else on_dev(noise.get("hard_mask")))
        noise = self.draw(B) if noise is None else noise
        drop, axis = noise["dropout"].to(dev), noise["axis"].to(dev).float()
        degree = math.pi * (torch.tanh(probs[:, :, 0:3]) * self.w_R_range) / 180.0 * drop[:, :, 0:1]
        scale = (torch.sigmoid(probs[:, :, 3:6]) * (self.w_S_range - 1) + 1) * drop[:, :, 1:2] * axis
        scale = scale + 1 * (scale == 0)
        trl = torch.tanh(probs[:, :, 6:9]) * self.w_T_range * drop[:, :, 2:3] * axis
        xyz_new = deform_mask(xyz, anchor, _affine(degree, scale, trl),
                              noise["proj_axis"].to(dev).float().reshape(B, 3), masking[:, :, 0], self.sigma)
        return xyz, xyz_new


# ----------------------------------------------------------------------------------------------------------
# discriminator
# ----------------------------------------------------------------------------------------------------------
class PointNetSetAbstraction_SpectralNorm(nn.Module):
    """group_all PointNet stage with spectral-norm 1x1 convs + ReLU, max over the points
    (point_discriminator.py:149-191; only group_all=True is used by the reference)."""

    def __init__(self, npoint, radius, nsample, in_channel, mlp, group_all):
        super().__init__()
        if not group_all:
            raise NotImplementedError("the AdaptPoint discriminator only uses group_all=True")
        self.mlp_convs = nn.ModuleList()
        self.mlp_bns = nn.ModuleList()
        last = in_channel
        for out in mlp:
            self.mlp_convs.append(spectral_norm(TcConv2d(last, out, 1)))
            last = out
        self.group_all = group_all

    def forward(self, xyz, points=None):
        """xyz (B,3,N) [, points (B,D,N)] -> (zeros (B,3,1), features (B,C,1))."""
        x = xyz if points is None else torch.cat([xyz, points], dim=1)
        x = x.unsqueeze(2)                                                    # (B,C,1,N) = permute of (B,1,N,C)
        for conv in self.mlp_convs[:-1]:
            x = F.relu(conv(x))
        # last layer: conv -> ReLU -> max over the points == ReLU(max over the points of the conv)
        pooled = F.relu(conv_max(self.mlp_convs[-1], x))                          # (B,C)
        return xyz.new_zeros(xyz.shape[0], xyz.shape[1], 1), pooled.unsqueeze(-1)


class PointDiscriminator1(nn.Module):
    def __init__(self, num_classes=40, normal_channel=False, **kwargs):
        super().__init__()
        sn = spectral_norm
        self.normal_channel = normal_channel
        self.sa1 = PointNetSetAbstraction_SpectralNorm(None, None, None, in_channel=3, mlp=[64, 128, 1024],
                                                       group_all=True)
        self.fc1 = sn(nn.Linear(1024, 512))
        self.drop1 = nn.Dropout(0.4)
        self.fc2 = sn(nn.Linear(512, 256))
        self.drop2 = nn.Dropout(0.4)
        self.fc3 = sn(nn.Linear(256, num_classes))
        self.prob_head = nn.Sequential(sn(nn.Linear(num_classes, 1)), nn.Sigmoid())

    def forward(self, xyz):
        """xyz (B,N,3) -> prob (B,1)."""
        xyz = xyz.permute(0, 2, 1).contiguous()
        B = xyz.shape[0]
        norm = xyz[:, 3:, :] if self.normal_channel else None
        _, feat = self.sa1(xyz[:, :3, :], norm)
        x = feat.view(B, 1024)
        x = self.drop1(F.relu(self.fc1(x)))
        x = self.drop2(F.relu(self.fc2(x)))
        return self.prob_head(self.fc3(x))
