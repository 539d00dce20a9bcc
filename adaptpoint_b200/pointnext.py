"""PointNeXt set-abstraction modules over adaptpoint_b200.ops (reference: openpoints/models/backbone/pointnext.py).

Same constructor arguments, forward semantics and parameter names as
  LocalAggregation  pointnext.py:27-78      SetAbstraction     pointnext.py:81-170
  FeaturePropogation pointnext.py:173-226   InvResMLP          pointnext.py:229-276
  PointNextEncoder  pointnext.py:310-456    PointNextDecoder   pointnext.py:459-496
so a reference state_dict (`encoder.{i}.{j}.convs.{l}.{0,1}`, `skipconv.0`, `pwconv.{l}.{0,1}`,
`decoder.{i}.0.convs...`) loads unchanged.  Configs are plain dicts (the reference's EasyDict/registry layer is
host infrastructure outside the path).

`fused=True` (SetAbstraction / LocalAggregation) routes group -> relative position -> shared MLP -> max-pool
through the tcgen05 kernels of csrc/sa_fused.cu whenever the block is in eval mode (BatchNorm folded) and autograd
is off; otherwise the reference's op sequence runs over the same C-ABI kernels.
"""
import copy
from typing import List

import torch
import torch.nn as nn

from . import geometry, ops
from .norm import ConvBlock, bn_max
from .layers import (CHANNEL_MAP, create_act, create_convblock1d, create_convblock2d, create_grouper,
                     get_aggregation_feautres, QueryAndGroup)


def _pool_fn(reduction):
    reduction = "mean" if reduction.lower() == "avg" else reduction.lower()
    if reduction == "max":
        return lambda x: torch.max(x, dim=-1, keepdim=False)[0]
    if reduction == "mean":
        return lambda x: torch.mean(x, dim=-1, keepdim=False)
    if reduction == "sum":
        return lambda x: torch.sum(x, dim=-1, keepdim=False)
    raise ValueError(reduction)


def _fusable(grouper, convs, feature_type, reduction, training):
    """The fused kernels (csrc/sa_fused.cu) cover: ball-query grouping with relative dp; 'dp_fj'; one or two
    (Conv2d 1x1 [+BN2d] [+ReLU]) layers; max reduction; BatchNorm in eval mode (folded); no autograd."""
    if training or torch.is_grad_enabled() or feature_type != "dp_fj" or reduction != "max":
        return False
    for blk in convs:
        if not isinstance(blk[0], nn.Conv2d) or any(not isinstance(m, (nn.Conv2d, nn.BatchNorm2d, nn.ReLU)) for m in blk):
            return False
    from . import fused
    return fused.fusable_reason(grouper, convs, getattr(grouper, "nsample", None)) is None


def _aggregate(mod, query_p, p, f):
    """grouper -> concat -> convs -> pool without the (B,C,M,K) tensors when a fused path covers the block:
      * eval mode, no autograd, `fused=True`: bf16 tcgen05 shared MLP (fused.py, csrc/sa_fused.cu);
      * one-conv blocks under autograd, training or eval mode (every block of PointNeXt-B/L/XL): fused_train.py,
        csrc/sa_train.cu -- forward and backward;
    otherwise the reference's op sequence over the same C-ABI kernels."""
    if mod.fused and _fusable(mod.grouper, mod.convs, mod.feature_type, mod.reduction, mod.training):
        from . import fused
        return fused.sa_fused_forward(mod, mod.grouper, mod.convs, query_p, p, f)
    from . import fused_train
    if fused_train.usable(mod, mod.grouper, mod.convs, mod.feature_type, mod.reduction, query_p, p, f):
        return fused_train.sa_one_layer(mod, mod.grouper, mod.convs, query_p, p, f)
    dp, fj = mod.grouper(query_p, p, f)
    return _convs_pool(mod, get_aggregation_feautres(query_p, dp, None, fj, "dp_fj"))


def _convs_pool(mod, x):
    """pool(convs(x)); when the pool is max and the last block is Conv -> BatchNorm [-> ReLU], the BatchNorm,
    activation and max over the neighbours run as norm.bn_max (one read pass, no normalised (B,C,M,K) tensor)."""
    last = mod.convs[-1]
    split = last.split_for_pool() if (isinstance(last, ConvBlock) and mod.reduction == "max") else None
    if split is None:
        return mod.pool(mod.convs(x))
    for blk in list(mod.convs)[:-1]:
        x = blk(x)
    conv, bn, relu = split
    return bn_max(conv(x), bn, relu)


class LocalAggregation(nn.Module):
    """Aggregate neighbour features around every point of the same set (pointnext.py:27-78)."""

    def __init__(self, channels: List[int], norm_args=None, act_args=None, group_args=None, conv_args=None,
                 feature_type="dp_fj", reduction="max", last_act=True, fused=False, **kwargs):
        super().__init__()
        norm_args = {"norm": "bn1d"} if norm_args is None else norm_args
        group_args = {"NAME": "ballquery", "radius": 0.1, "nsample": 16} if group_args is None else group_args
        channels = list(channels)
        channels[0] = CHANNEL_MAP[feature_type](channels[0])
        n = len(channels) - 1
        self.convs = nn.Sequential(*[
            create_convblock2d(channels[i], channels[i + 1], norm_args=norm_args,
                               act_args=None if (i == n - 1 and not last_act) else act_args, **(conv_args or {}))
            for i in range(n)])
        self.grouper = create_grouper(group_args)
        self.reduction = "mean" if reduction.lower() == "avg" else reduction.lower()
        self.pool = _pool_fn(reduction)
        self.feature_type = feature_type
        self.fused = fused

    def forward(self, pf):
        p, f = pf
        if self.feature_type == "dp_fj":
            return _aggregate(self, p, p, f)
        dp, fj = self.grouper(p, p, f)
        return _convs_pool(self, get_aggregation_feautres(p, dp, f, fj, self.feature_type))


class SetAbstraction(nn.Module):
    """Sub-sample (FPS), group, shared MLP, max-pool, optional residual (pointnext.py:81-170)."""

    def __init__(self, in_channels, out_channels, layers=1, stride=1, group_args=None, norm_args=None,
                 act_args=None, conv_args=None, sampler="fps", feature_type="dp_fj", use_res=False, is_head=False,
                 fused=False, **kwargs):
        super().__init__()
        norm_args = {"norm": "bn1d"} if norm_args is None else norm_args
        act_args = {"act": "relu"} if act_args is None else act_args
        group_args = copy.deepcopy(dict(group_args or {"NAME": "ballquery", "radius": 0.1, "nsample": 16}))
        self.stride, self.is_head = stride, is_head
        self.all_aggr = not is_head and stride == 1
        self.use_res = use_res and not self.all_aggr and not is_head
        self.feature_type = feature_type
        self.reduction = "max"
        self.fused = fused

        mid = out_channels // 2 if stride > 1 else out_channels
        channels = [in_channels] + [mid] * (layers - 1) + [out_channels]
        if not is_head:
            channels[0] = CHANNEL_MAP[feature_type](channels[0])
        if self.use_res:
            self.skipconv = (create_convblock1d(in_channels, channels[-1], norm_args=None, act_args=None)
                             if in_channels != channels[-1] else nn.Identity())
            self.act = create_act(act_args)
        make = create_convblock1d if is_head else create_convblock2d
        n = len(channels) - 1
        self.convs = nn.Sequential(*[
            make(channels[i], channels[i + 1], norm_args=None if is_head else norm_args,
                 act_args=None if (i == n - 1 and (self.use_res or is_head)) else act_args, **(conv_args or {}))
            for i in range(n)])
        if not is_head:
            if self.all_aggr:
                group_args["nsample"] = None
                group_args["radius"] = None
            self.grouper = create_grouper(group_args)
            self.pool = _pool_fn("max")
            if sampler.lower() != "fps":
                raise NotImplementedError("only the 'fps' sampler is on the accelerated path")
            self.sample_fn = ops.furthest_point_sample

    def forward(self, pf):
        p, f = pf
        if self.is_head:
            return p, self.convs(f)
        if self.all_aggr:
            dp, fj = self.grouper(p, p, f)
            return p, _convs_pool(self, get_aggregation_feautres(p, dp, None, fj, self.feature_type))
        if self.sample_fn is ops.furthest_point_sample:
            idx, new_p = geometry.sample(p, p.shape[1] // self.stride)       # shared FPS chain when p belongs to a plan
        else:
            idx = self.sample_fn(p, p.shape[1] // self.stride).long()
            new_p = torch.gather(p, 1, idx.unsqueeze(-1).expand(-1, -1, 3))
        fi = identity = None
        if self.use_res or "df" in self.feature_type:
            fi = torch.gather(f, -1, idx.unsqueeze(1).expand(-1, f.shape[1], -1))
            if self.use_res:
                identity = self.skipconv(fi)
        if self.feature_type == "dp_fj":
            out = _aggregate(self, new_p, p, f)
        else:
            dp, fj = self.grouper(new_p, p, f)
            out = _convs_pool(self, get_aggregation_feautres(new_p, dp, fi, fj, self.feature_type))
        if self.use_res:
            out = self.act(out + identity)
        return new_p, out


class FeaturePropogation(nn.Module):
    """PointNet++ feature propagation: 3-NN inverse-distance interpolation + skip concat + MLP
    (pointnext.py:173-226; the reference's spelling of the class name is kept)."""

    def __init__(self, mlp, upsample=True, norm_args=None, act_args=None):
        super().__init__()
        norm_args = {"norm": "bn1d"} if norm_args is None else norm_args
        act_args = {"act": "relu"} if act_args is None else act_args
        mlp = list(mlp)
        if not upsample:
            self.linear2 = nn.Sequential(nn.Linear(mlp[0], mlp[1]), nn.ReLU(inplace=True))
            mlp[1] *= 2
            self.linear1 = nn.Sequential(*[create_convblock1d(mlp[i], mlp[i + 1], norm_args=norm_args,
                                                              act_args=act_args) for i in range(1, len(mlp) - 1)])
        else:
            self.convs = nn.Sequential(*[create_convblock1d(mlp[i], mlp[i + 1], norm_args=norm_args,
                                                            act_args=act_args) for i in range(len(mlp) - 1)])

    def forward(self, pf1, pf2=None):
        if pf2 is None:
            _, f = pf1
            g = self.linear2(f.mean(dim=-1))
            return self.linear1(torch.cat((f, g.unsqueeze(-1).expand(-1, -1, f.shape[-1])), dim=1))
        p1, f1 = pf1
        p2, f2 = pf2
        up = ops.three_interpolation(p1, p2, f2)
        return self.convs(up if f1 is None else torch.cat((f1, up), dim=1))


class InvResMLP(nn.Module):
    """LocalAggregation + point-wise inverted-bottleneck MLP + residual (pointnext.py:229-276)."""

    def __init__(self, in_channels, norm_args=None, act_args=None, aggr_args=None, group_args=None,
                 conv_args=None, expansion=1, use_res=True, num_posconvs=2, less_act=False, fused=False, **kwargs):
        super().__init__()
        aggr_args = {"feature_type": "dp_fj", "reduction": "max"} if aggr_args is None else aggr_args
        group_args = {"NAME": "ballquery"} if group_args is None else group_args
        self.use_res = use_res
        mid = int(in_channels * expansion)
        self.convs = LocalAggregation([in_channels, in_channels], norm_args=norm_args,
                                      act_args=act_args if num_posconvs > 0 else None, group_args=group_args,
                                      conv_args=conv_args, fused=fused, **aggr_args, **kwargs)
        channels = [] if num_posconvs < 1 else ([in_channels, in_channels] if num_posconvs == 1
                                                else [in_channels, mid, in_channels])
        n = len(channels) - 1
        self.pwconv = nn.Sequential(*[
            create_convblock1d(channels[i], channels[i + 1], norm_args=norm_args,
                               act_args=act_args if (i != n - 1 and not less_act) else None, **(conv_args or {}))
            for i in range(n)])
        self.act = create_act(act_args)

    def forward(self, pf):
        p, f = pf
        identity = f
        f = self.pwconv(self.convs([p, f]))
        if f.shape[-1] == identity.shape[-1] and self.use_res:
            f = f + identity
        return [p, self.act(f)]


class PointNextEncoder(nn.Module):
    """Stage wiring and the radius / nsample schedule (pointnext.py:310-456)."""

    def __init__(self, in_channels=4, width=32, blocks=(1, 4, 7, 4, 4), strides=(4, 4, 4, 4), block="InvResMLP",
                 nsample=32, radius=0.1, aggr_args=None, group_args=None, sa_layers=1, sa_use_res=False,
                 fused=False, **kwargs):
        super().__init__()
        if block not in ("InvResMLP", InvResMLP):
            raise NotImplementedError("only InvResMLP blocks are on the accelerated path")
        self.blocks, self.strides = list(blocks), list(strides)
        self.in_channels = in_channels
        self.aggr_args = aggr_args or {"feature_type": "dp_fj", "reduction": "max"}
        self.norm_args = kwargs.get("norm_args", {"norm": "bn"})
        self.act_args = kwargs.get("act_args", {"act": "relu"})
        self.conv_args = kwargs.get("conv_args", None)
        self.sampler = kwargs.get("sampler", "fps")
        self.expansion = kwargs.get("expansion", 4)
        self.sa_layers, self.sa_use_res = sa_layers, sa_use_res
        self.use_res = kwargs.get("use_res", True)
        self.fused = fused
        self.radii = self._to_full_list(radius, kwargs.get("radius_scaling", 2))
        self.nsample = self._to_full_list(nsample, kwargs.get("nsample_scaling", 1))
        group_args = dict(group_args or {"NAME": "ballquery"})

        channels = []
        for s in self.strides:
            if s != 1:
                width *= 2
            channels.append(width)
        stages = []
        for i in range(len(self.blocks)):
            stages.append(self._make_enc(channels[i], self.blocks[i], self.strides[i], group_args, self.radii[i],
                                         self.nsample[i], is_head=(i == 0 and self.strides[i] == 1)))
        self.encoder = nn.Sequential(*stages)
        self.out_channels = channels[-1]
        self.channel_list = channels

    def _to_full_list(self, param, scaling=1):
        """Per-stage, per-block value lists (pointnext.py:389-407)."""
        full = []
        if isinstance(param, (list, tuple)):
            for i, value in enumerate(param):
                value = list(value) if isinstance(value, (list, tuple)) else [value]
                value = value + [value[-1]] * (self.blocks[i] - len(value))
                full.append(value)
            return full
        for i, s in enumerate(self.strides):
            if s == 1:
                full.append([param] * self.blocks[i])
            else:
                full.append([param] + [param * scaling] * (self.blocks[i] - 1))
                param *= scaling
        return full

    def _make_enc(self, channels, nblocks, stride, group_args, radii, nsamples, is_head):
        ga = dict(group_args, radius=radii[0], nsample=nsamples[0])
        layers = [SetAbstraction(self.in_channels, channels, self.sa_layers if not is_head else 1, stride,
                                 group_args=ga, sampler=self.sampler, norm_args=self.norm_args,
                                 act_args=self.act_args, conv_args=self.conv_args, is_head=is_head,
                                 use_res=self.sa_use_res, fused=self.fused, **self.aggr_args)]
        self.in_channels = channels
        for i in range(1, nblocks):
            ga = dict(group_args, radius=radii[i], nsample=nsamples[i])
            layers.append(InvResMLP(channels, aggr_args=self.aggr_args, norm_args=self.norm_args,
                                    act_args=self.act_args, group_args=ga, conv_args=self.conv_args,
                                    expansion=self.expansion, use_res=self.use_res, fused=self.fused))
        return nn.Sequential(*layers)

    def geometry_spec(self, n):
        """(chains, queries) of one forward over n-point clouds, in geometry.SamplingPlan / GeometryMailbox terms: the
        FPS chain of the strided stages, the ball query of every stage-opening SetAbstraction (support = parent level)
        and the one query the InvResMLP blocks of a stage share (the level querying itself)."""
        chain, queries, path = [], [], ()
        for stage in self.encoder:
            sa = stage[0]
            if getattr(sa, "stride", 1) > 1 and not sa.is_head:
                n = n // sa.stride
                child = path + (n,)
                chain.append(n)
                g = sa.grouper
                if getattr(g, "radius", None) is not None and getattr(g, "nsample", None) is not None:
                    queries.append((path, child, float(g.radius), int(g.nsample)))
                path = child
            for blk in list(stage)[1:]:
                g = getattr(getattr(blk, "convs", None), "grouper", None)
                if getattr(g, "radius", None) is not None and getattr(g, "nsample", None) is not None:
                    q = (path, path, float(g.radius), int(g.nsample))
                    if q not in queries:
                        queries.append(q)
        return [chain], queries

    @staticmethod
    def _unpack(p0, f0):
        if hasattr(p0, "keys"):
            p0, f0 = p0["pos"], p0.get("x", None)
        if hasattr(p0, "__dict__"):
            p0.__dict__.pop("_apb_bq", None)     # the ball-query memo (geometry.ball_query) lives for ONE forward
        from . import fused_train
        fused_train.reset_step_cache()
        if f0 is None:
            f0 = p0.clone().transpose(1, 2).contiguous()
        return p0, f0

    def forward_cls_feat(self, p0, f0=None):
        p0, f0 = self._unpack(p0, f0)
        for stage in self.encoder:
            p0, f0 = stage([p0, f0])
        return f0.squeeze(-1)

    def forward_seg_feat(self, p0, f0=None):
        p0, f0 = self._unpack(p0, f0)
        p, f = [p0], [f0]
        for stage in self.encoder:
            _p, _f = stage([p[-1], f[-1]])
            p.append(_p)
            f.append(_f)
        return p, f

    def forward(self, p0, f0=None):
        return self.forward_seg_feat(p0, f0)


class PointNextDecoder(nn.Module):
    """Feature-propagation decoder, deepest stage first (pointnext.py:459-496)."""

    def __init__(self, encoder_channel_list: List[int], decoder_layers=2, decoder_stages=4, **kwargs):
        super().__init__()
        self.decoder_layers = decoder_layers
        self.in_channels = encoder_channel_list[-1]
        skip = list(encoder_channel_list[:-1])
        if len(skip) < decoder_stages:
            skip.insert(0, kwargs.get("in_channels", 3))
        fp = list(encoder_channel_list[:decoder_stages])
        n = len(fp)
        stages = [None] * n
        for i in range(-1, -n - 1, -1):
            mlp = [skip[i] + self.in_channels] + [fp[i]] * decoder_layers
            stages[i] = nn.Sequential(FeaturePropogation(mlp))
            self.in_channels = fp[i]
        self.decoder = nn.Sequential(*stages)
        self.out_channels = fp[-n]

    def forward(self, p, f):
        for i in range(-1, -len(self.decoder) - 1, -1):
            f[i - 1] = self.decoder[i][0]([p[i - 1], f[i - 1]], [p[i], f[i]])
        return f[-len(self.decoder) - 1]
