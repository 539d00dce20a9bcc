"""Layer factories and neighbourhood groupers of the set-abstraction path, over adaptpoint_b200.ops.

Mirrors (names, argument meaning, resulting nn.Sequential layout -- so reference checkpoints load):
  openpoints/models/layers/conv.py:24-146        create_convblock1d / 2d / create_linearblock
  openpoints/models/layers/norm.py:57-98         create_norm   ('bn' + dimension -> BatchNorm{1,2}d)
  openpoints/models/layers/activation.py:5-52    create_act
  openpoints/models/layers/group.py:12-28,206-352  KNN, QueryAndGroup, GroupAll, KNNGroup,
                                                   get_aggregation_feautres, create_grouper
  openpoints/models/layers/local_aggregation.py:13-29  CHANNEL_MAP
Only the 1x1-conv / BatchNorm / pointwise-activation subset that PointNeXt and the AdaptPoint imitator
use is provided; anything else raises (no silent substitution).
"""
import copy

import torch
import torch.nn as nn

from . import geometry, ops
from .norm import ConvBlock, TcBatchNorm1d, TcBatchNorm2d
from .tcconv import TcConv1d, TcConv2d

_ACTS = {
    "relu": nn.ReLU, "leakyrelu": nn.LeakyReLU, "leaky_relu": nn.LeakyReLU, "gelu": nn.GELU,
    "silu": nn.SiLU, "swish": nn.SiLU, "sigmoid": nn.Sigmoid, "tanh": nn.Tanh, "elu": nn.ELU,
}
_NORMS = {"bn1d": TcBatchNorm1d, "bn2d": TcBatchNorm2d, "syncbn": nn.SyncBatchNorm,
          "in1d": nn.InstanceNorm1d, "in2d": nn.InstanceNorm2d}


def create_act(act_args):
    """`{'act': name, ...}` or a name -> module (inplace by default, like the reference)."""
    if act_args is None:
        return None
    cfg = {"act": act_args} if isinstance(act_args, str) else dict(act_args)
    name = cfg.pop("act", None)
    if name is None:
        return None
    name = name.lower()
    if name not in _ACTS:
        raise NotImplementedError(f"activation {name!r}")
    inplace = cfg.pop("inplace", True)
    if name in ("gelu", "sigmoid", "tanh"):
        return _ACTS[name](**cfg)
    return _ACTS[name](inplace=inplace, **cfg)


def create_norm(norm_args, channels, dimension=None):
    """`{'norm': 'bn'}` + dimension '1d'/'2d' -> BatchNorm1d/2d (norm.py:75-98)."""
    if norm_args is None:
        return None
    cfg = dict(norm_args) if isinstance(norm_args, dict) else {"norm": norm_args}
    name = cfg.pop("norm", None)
    if name is None:
        return None
    name = name.lower()
    if dimension is not None and str(dimension).lower() not in name:
        name += str(dimension).lower()
    if name not in _NORMS:
        raise NotImplementedError(f"norm {name!r}")
    return _NORMS[name](channels, **cfg)


def _block(make_linear, dim, cin, cout, norm_args, act_args, order, bias):
    if order == "conv-norm-act":
        norm = create_norm(norm_args, cout, dim)
        mods = [make_linear(cin, cout, bias and norm is None)]
        if norm is not None:
            mods.append(norm)
        if act_args is not None:
            mods.append(create_act(act_args))
    elif order == "norm-act-conv":
        norm = create_norm(norm_args, cin, dim)
        mods = [] if norm is None else [norm]
        if act_args is not None:
            mods.append(create_act(act_args))
        mods.append(make_linear(cin, cout, bias and norm is None))
    elif order == "conv-act-norm":
        norm = create_norm(norm_args, cout, dim)
        mods = [make_linear(cin, cout, bias and norm is None)]
        if act_args is not None:
            mods.append(create_act(act_args))
        if norm is not None:
            mods.append(norm)
    else:
        raise NotImplementedError(f"{order} is not supported")
    return ConvBlock(*mods)


def create_convblock2d(cin, cout, norm_args=None, act_args=None, order="conv-norm-act", bias=True, **kw):
    """1x1 Conv2d (+norm)(+act); conv bias is dropped when a norm follows (conv.py:24-60)."""
    return _block(lambda i, o, b: TcConv2d(i, o, (1, 1), bias=b, **kw), "2d", cin, cout, norm_args, act_args,
                  order, bias)


def create_convblock1d(cin, cout, norm_args=None, act_args=None, order="conv-norm-act", bias=True, **kw):
    """1x1 Conv1d (+norm)(+act) (conv.py:63-103)."""
    return _block(lambda i, o, b: TcConv1d(i, o, 1, bias=b, **kw), "1d", cin, cout, norm_args, act_args, order,
                  bias)


def create_linearblock(cin, cout, norm_args=None, act_args=None, order="conv-norm-act", bias=True, **kw):
    """Linear (+BatchNorm1d)(+act) (conv.py:106-146)."""
    return _block(lambda i, o, b: nn.Linear(i, o, b, **kw), "1d", cin, cout, norm_args, act_args, order, bias)


CHANNEL_MAP = {
    "fj": lambda c: c, "df": lambda c: c, "dp": lambda c: 3, "dp_fj": lambda c: 3 + c, "dp_df": lambda c: c + 3,
    "dp_fj_df": lambda c: 2 * c + 3, "pi_dp_fj_df": lambda c: 2 * c + 6,
}


def get_aggregation_feautres(p, dp, f, fj, feature_type="dp_fj"):
    """Channel concat of the grouped tensors (group.py:323-335; the reference's spelling is kept)."""
    if feature_type == "dp_fj":
        return torch.cat([dp, fj], 1)
    df = fj - f.unsqueeze(-1)
    if feature_type == "dp_fj_df":
        return torch.cat([dp, fj, df], 1)
    if feature_type == "pi_dp_fj_df":
        pi = p.transpose(1, 2).unsqueeze(-1).expand(-1, -1, -1, df.shape[-1])
        return torch.cat([pi, dp, fj, df], 1)
    if feature_type == "dp_df":
        return torch.cat([dp, df], 1)
    raise NotImplementedError(feature_type)


class KNN(nn.Module):
    """k nearest support points per query (group.py:12-28).  The reference materialises
    cdist(support, query) and calls topk; here apb_knn_f32 streams the support cloud through shared
    memory.  Returns (distances (B,K,M) like topk(dim=1).values, idx (B,M,K) int32)."""

    def __init__(self, neighbors, transpose_mode=True):
        super().__init__()
        self.neighbors = neighbors

    @torch.no_grad()
    def forward(self, support, query):
        idx, d2 = ops.knn_query(self.neighbors, support.contiguous(), query.contiguous(), return_dist2=True)
        return d2.sqrt().transpose(1, 2), idx


class QueryAndGroup(nn.Module):
    """Ball query + grouping (group.py:206-255)."""

    def __init__(self, radius, nsample, relative_xyz=True, normalize_dp=False, return_only_idx=False, **kwargs):
        super().__init__()
        self.radius, self.nsample = radius, nsample
        self.relative_xyz, self.normalize_dp, self.return_only_idx = relative_xyz, normalize_dp, return_only_idx

    def forward(self, query_xyz, support_xyz, features=None):
        idx = geometry.ball_query(self.radius, self.nsample, support_xyz, query_xyz)
        if self.return_only_idx:
            return idx
        grouped_xyz = ops.grouping_operation(support_xyz.transpose(1, 2).contiguous(), idx)
        if self.relative_xyz:
            grouped_xyz = grouped_xyz - query_xyz.transpose(1, 2).unsqueeze(-1)
            if self.normalize_dp:
                grouped_xyz = grouped_xyz / self.radius
        grouped_features = ops.grouping_operation(features, idx) if features is not None else None
        return grouped_xyz, grouped_features


class GroupAll(nn.Module):
    """One group holding every point; `dp` is the ABSOLUTE xyz (group.py:258-272)."""

    def forward(self, new_xyz, xyz, features=None):
        return xyz.transpose(1, 2).unsqueeze(2), (features.unsqueeze(2) if features is not None else None)


class KNNGroup(nn.Module):
    """kNN + grouping; `normalize_dp` divides by the per-cloud max neighbour distance (group.py:275-320)."""

    def __init__(self, nsample, relative_xyz=True, normalize_dp=False, return_only_idx=False, **kwargs):
        super().__init__()
        self.nsample = nsample
        self.knn = KNN(nsample)
        self.relative_xyz, self.normalize_dp, self.return_only_idx = relative_xyz, normalize_dp, return_only_idx

    def forward(self, query_xyz, support_xyz, features=None):
        _, idx = self.knn(support_xyz, query_xyz)
        if self.return_only_idx:
            return idx
        grouped_xyz = ops.grouping_operation(support_xyz.transpose(1, 2).contiguous(), idx)
        if self.relative_xyz:
            grouped_xyz = grouped_xyz - query_xyz.transpose(1, 2).unsqueeze(-1)
        if self.normalize_dp:
            grouped_xyz = grouped_xyz / torch.amax(grouped_xyz.pow(2).sum(1).sqrt(), dim=(1, 2)).view(-1, 1, 1, 1)
        return grouped_xyz, (ops.grouping_operation(features, idx) if features is not None else None)


def create_grouper(group_args):
    """`{'NAME': 'ballquery'|'knn', 'radius', 'nsample' (None -> GroupAll), ...}` (group.py:338-352)."""
    cfg = copy.deepcopy(dict(group_args))
    method = cfg.pop("NAME", "ballquery")
    radius = cfg.pop("radius", 0.1)
    nsample = cfg.pop("nsample", 20)
    if nsample is None:
        return GroupAll()
    if method == "ballquery":
        return QueryAndGroup(radius, nsample, **cfg)
    if method == "knn":
        return KNNGroup(nsample, **cfg)
    raise NotImplementedError(method)
