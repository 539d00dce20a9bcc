"""Classification / segmentation wrappers and the model configurations of BASELINE.json.

  BaseCls, ClsHead        openpoints/models/classification/cls_base.py:12-39, 76-136
  BaseSeg, SegHead        openpoints/models/segmentation/base_seg.py:14-50, 88-141
  SmoothCrossEntropy      openpoints/loss/build.py:11-63
Parameter names match the reference (`encoder.encoder.*`, `prediction.head.*`, `decoder.decoder.*`, `head.head.*`).

CONFIGS: 'pointnext-s' is cfgs/scanobjectnn/pointnext-s.yaml:5-35.  The reference ships no B/L/XL YAML
(SURVEY.md 2.1 #28); those entries are authored from the encoder's constructor defaults
(pointnext.py:335-365) plus the upstream PointNeXt scaling table: width 32 with blocks [1,2,3,2,2] (B) /
[1,3,5,3,3] (L), width 64 with blocks [1,4,7,4,4] (XL), strides [1,4,4,4,4], radius 0.1 x2 per stage, nsample 32.
"""
import copy

import torch
import torch.nn as nn
import torch.nn.functional as F

from .layers import create_convblock1d, create_linearblock
from .pointnext import PointNextDecoder, PointNextEncoder


class SmoothCrossEntropy(nn.Module):
    """Label-smoothed cross entropy: target = (1-s) on the label, s/(C-1) elsewhere (loss/build.py:37-63)."""

    def __init__(self, label_smoothing=0.2, **kwargs):
        super().__init__()
        self.label_smoothing = label_smoothing

    def forward(self, pred, gt):
        if pred.dim() > 2:
            pred = pred.transpose(1, 2).reshape(-1, pred.shape[1])
        gt = gt.contiguous().view(-1)
        if self.label_smoothing <= 0:
            return F.cross_entropy(pred, gt)
        n_class = pred.size(1)
        one_hot = torch.zeros_like(pred).scatter(1, gt.view(-1, 1), 1)
        target = one_hot * (1 - self.label_smoothing) + (1 - one_hot) * self.label_smoothing / (n_class - 1)
        return -(target * F.log_softmax(pred, dim=1)).sum(dim=1).mean()


class ClsHead(nn.Module):
    """MLP classification head with optional global pooling of point features (cls_base.py:76-136)."""

    def __init__(self, num_classes, in_channels, mlps=(256,), norm_args=None, act_args=None, dropout=0.5,
                 global_feat=None, point_dim=2, **kwargs):
        super().__init__()
        act_args = {"act": "relu"} if act_args is None else act_args
        self.global_feat = global_feat.split(",") if global_feat is not None else None
        self.point_dim = point_dim
        if self.global_feat is not None:
            in_channels *= len(self.global_feat)
        dims = [in_channels] + list(mlps or []) + [num_classes]
        heads = []
        for i in range(len(dims) - 2):
            heads.append(create_linearblock(dims[i], dims[i + 1], norm_args=norm_args, act_args=act_args))
            if dropout:
                heads.append(nn.Dropout(dropout))
        heads.append(create_linearblock(dims[-2], dims[-1], act_args=None))
        self.head = nn.Sequential(*heads)

    def forward(self, end_points):
        if self.global_feat is not None:
            feats = []
            for kind in self.global_feat:
                if "max" in kind:
                    feats.append(torch.max(end_points, dim=self.point_dim, keepdim=False)[0])
                elif kind in ("avg", "mean"):
                    feats.append(torch.mean(end_points, dim=self.point_dim, keepdim=False))
            end_points = torch.cat(feats, dim=1)
        return self.head(end_points)


class BaseCls(nn.Module):
    """encoder.forward_cls_feat -> prediction head (cls_base.py:12-39)."""

    def __init__(self, encoder_args=None, cls_args=None, criterion_args=None, **kwargs):
        super().__init__()
        enc = copy.deepcopy(dict(encoder_args))
        enc.pop("NAME", None)
        self.encoder = PointNextEncoder(**enc)
        if cls_args is not None:
            cls = copy.deepcopy(dict(cls_args))
            cls.pop("NAME", None)
            cls["in_channels"] = self.encoder.out_channels
            self.prediction = ClsHead(**cls)
        else:
            self.prediction = nn.Identity()
        if criterion_args is not None:
            crit = dict(criterion_args)
            crit.pop("NAME", None)
            self.criterion = SmoothCrossEntropy(**crit)
        else:
            self.criterion = None

    def forward(self, data):
        return self.prediction(self.encoder.forward_cls_feat(data))

    def get_loss(self, pred, gt, inputs=None):
        return self.criterion(pred, gt.long())

    def get_logits_loss(self, data, gt):
        logits = self.forward(data)
        return logits, self.criterion(logits, gt.long())


class SegHead(nn.Module):
    """Per-point Conv1d head, optional global-feature concat (base_seg.py:88-141)."""

    def __init__(self, num_classes, in_channels, mlps=None, norm_args=None, act_args=None, dropout=0.5,
                 global_feat=None, **kwargs):
        super().__init__()
        norm_args = {"norm": "bn1d"} if norm_args is None else norm_args
        act_args = {"act": "relu"} if act_args is None else act_args
        self.global_feat = global_feat.split(",") if global_feat is not None else None
        in_channels *= (len(self.global_feat) + 1) if self.global_feat is not None else 1
        if mlps is None:
            dims = [in_channels, in_channels, num_classes]
        else:
            dims = [in_channels] + (list(mlps) if isinstance(mlps, (list, tuple)) else [mlps]) + [num_classes]
        heads = []
        for i in range(len(dims) - 2):
            heads.append(create_convblock1d(dims[i], dims[i + 1], norm_args=norm_args, act_args=act_args))
            if dropout:
                heads.append(nn.Dropout(dropout))
        heads.append(create_convblock1d(dims[-2], dims[-1], act_args=None))
        self.head = nn.Sequential(*heads)

    def forward(self, end_points):
        if self.global_feat is not None:
            feats = []
            for kind in self.global_feat:
                if "max" in kind:
                    feats.append(torch.max(end_points, dim=-1, keepdim=True)[0])
                elif kind in ("avg", "mean"):
                    feats.append(torch.mean(end_points, dim=-1, keepdim=True))
            g = torch.cat(feats, dim=1).expand(-1, -1, end_points.shape[-1])
            end_points = torch.cat((end_points, g), dim=1)
        return self.head(end_points)


class BaseSeg(nn.Module):
    """encoder.forward_seg_feat -> PointNextDecoder -> SegHead (base_seg.py:14-50)."""

    def __init__(self, encoder_args=None, decoder_args=None, cls_args=None, **kwargs):
        super().__init__()
        enc = copy.deepcopy(dict(encoder_args))
        enc.pop("NAME", None)
        self.encoder = PointNextEncoder(**enc)
        if decoder_args is not None:
            dec = copy.deepcopy(enc)
            dec.update({k: v for k, v in dict(decoder_args).items() if k != "NAME"})
            dec["encoder_channel_list"] = list(self.encoder.channel_list)
            self.decoder = PointNextDecoder(**dec)
        else:
            self.decoder = None
        if cls_args is not None:
            cls = copy.deepcopy(dict(cls_args))
            cls.pop("NAME", None)
            cls["in_channels"] = self.decoder.out_channels if self.decoder is not None else self.encoder.out_channels
            self.head = SegHead(**cls)
        else:
            self.head = None

    def forward(self, data):
        p, f = self.encoder.forward_seg_feat(data)
        if self.decoder is not None:
            f = self.decoder(p, f).squeeze(-1)
        if self.head is not None:
            f = self.head(f)
        return f


def _encoder(width, blocks, in_channels=4, **over):
    enc = dict(NAME="PointNextEncoder", blocks=blocks, strides=[1, 4, 4, 4, 4], width=width,
               in_channels=in_channels, sa_layers=1, sa_use_res=False, radius=0.1, radius_scaling=2, nsample=32,
               expansion=4, aggr_args={"feature_type": "dp_fj", "reduction": "max"},
               group_args={"NAME": "ballquery", "normalize_dp": True}, conv_args={"order": "conv-norm-act"},
               act_args={"act": "relu"}, norm_args={"norm": "bn"})
    enc.update(over)
    return enc


CONFIGS = {
    # cfgs/scanobjectnn/pointnext-s.yaml:5-35 + criterion of cfgs/scanobjectnn/default.yaml:35-37
    "pointnext-s": dict(
        kind="cls",
        encoder_args=_encoder(32, [1, 1, 1, 1, 1, 1], strides=[1, 2, 2, 2, 2, 1], sa_layers=2, sa_use_res=True,
                              radius=0.15, radius_scaling=1.5),
        cls_args=dict(NAME="ClsHead", num_classes=15, mlps=[512, 256], norm_args={"norm": "bn1d"}),
        criterion_args=dict(NAME="SmoothCrossEntropy", label_smoothing=0.3)),
    # ModelNet40-shaped classification (BASELINE.json config 4); stride-4 stages end with 8192/256 = 32 points,
    # pooled by ClsHead(global_feat='max') as cls_base.py:107-134 provides.
    "pointnext-b-cls": dict(
        kind="cls", encoder_args=_encoder(32, [1, 2, 3, 2, 2]),
        cls_args=dict(NAME="ClsHead", num_classes=40, mlps=[512, 256], norm_args={"norm": "bn1d"},
                      global_feat="max"),
        criterion_args=dict(NAME="SmoothCrossEntropy", label_smoothing=0.2)),
    "pointnext-l-cls": dict(
        kind="cls", encoder_args=_encoder(32, [1, 3, 5, 3, 3]),
        cls_args=dict(NAME="ClsHead", num_classes=40, mlps=[512, 256], norm_args={"norm": "bn1d"},
                      global_feat="max"),
        criterion_args=dict(NAME="SmoothCrossEntropy", label_smoothing=0.2)),
    # S3DIS-shaped segmentation (BASELINE.json config 5); params should land near docs/modelzoo.md:29-32.
    "pointnext-s-seg": dict(kind="seg", encoder_args=_encoder(32, [1, 1, 1, 1, 1]),
                            decoder_args=dict(NAME="PointNextDecoder"),
                            cls_args=dict(NAME="SegHead", num_classes=13)),
    "pointnext-b-seg": dict(kind="seg", encoder_args=_encoder(32, [1, 2, 3, 2, 2]),
                            decoder_args=dict(NAME="PointNextDecoder"),
                            cls_args=dict(NAME="SegHead", num_classes=13)),
    "pointnext-l-seg": dict(kind="seg", encoder_args=_encoder(32, [1, 3, 5, 3, 3]),
                            decoder_args=dict(NAME="PointNextDecoder"),
                            cls_args=dict(NAME="SegHead", num_classes=13)),
    "pointnext-xl-seg": dict(kind="seg", encoder_args=_encoder(64, [1, 4, 7, 4, 4]),
                             decoder_args=dict(NAME="PointNextDecoder"),
                             cls_args=dict(NAME="SegHead", num_classes=13)),
}


def build_model(name_or_cfg, fused=False, **encoder_over):
    """Instantiate a CONFIGS entry (or a cfg dict of the same shape)."""
    cfg = copy.deepcopy(CONFIGS[name_or_cfg] if isinstance(name_or_cfg, str) else name_or_cfg)
    cfg["encoder_args"].update(encoder_over)
    cfg["encoder_args"]["fused"] = fused
    kind = cfg.pop("kind", "cls")
    return BaseCls(**cfg) if kind == "cls" else BaseSeg(**cfg)
