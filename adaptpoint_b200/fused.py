"""Host side of the tensor-core shared-MLP kernels (csrc/sa_fused.cu): BatchNorm folding, bf16 weight packing and
the two calls `sa_fused_forward` / `pwconv_forward`.

They stand in for these reference op chains when the block is in eval mode (BatchNorm uses running statistics):
  SetAbstraction.forward      pointnext.py:164-166   grouper -> cat([dp, fj]) -> convs -> max over K
  LocalAggregation.forward    pointnext.py:66-77     same, query set == support set
  create_convblock1d blocks   conv.py:63-103         Conv1d 1x1 (+BN)(+ReLU)
Inference only: the outputs carry no autograd graph (the training path is the unfused op sequence over the same
C-ABI kernels).  Numerics: bf16 operands, fp32 accumulation -> compare at 2e-2 relative (BASELINE.json north_star).
"""
import torch
import torch.nn as nn

from . import _cabi, geometry, ops

_BK = 64  # reduction chunk of the kernel: weights are zero-padded to a multiple of this


def fold_conv_block(block):
    """nn.Sequential(Conv{1,2}d 1x1, [BatchNorm], [ReLU]) -> (W (Cout,Cin) f32, scale (Cout), shift (Cout), relu).
    Eval-mode BatchNorm: y = scale * (W x) + shift with scale = gamma / sqrt(var + eps),
    shift = beta - mean * scale (+ scale * conv_bias)."""
    mods = list(block)
    conv = mods[0]
    if not isinstance(conv, (nn.Conv1d, nn.Conv2d)) or any(k != 1 for k in conv.kernel_size) or conv.groups != 1:
        raise NotImplementedError("fused path covers 1x1, groups=1 convolutions only")
    W = conv.weight.detach().reshape(conv.out_channels, conv.in_channels).float()
    dev = W.device
    scale = torch.ones(conv.out_channels, device=dev)
    shift = torch.zeros(conv.out_channels, device=dev)
    if conv.bias is not None:
        shift = conv.bias.detach().float().clone()
    relu = False
    for m in mods[1:]:
        if isinstance(m, (nn.BatchNorm1d, nn.BatchNorm2d, nn.SyncBatchNorm)):
            if relu:
                raise NotImplementedError("norm after activation")
            g = m.weight.detach().float() if m.affine else torch.ones_like(scale)
            b = m.bias.detach().float() if m.affine else torch.zeros_like(scale)
            s = g / torch.sqrt(m.running_var.detach().float() + m.eps)
            shift = b + (shift - m.running_mean.detach().float()) * s
            scale = scale * s
        elif isinstance(m, nn.ReLU):
            relu = True
        else:
            raise NotImplementedError(f"fused path: unsupported layer {type(m).__name__}")
    return W, scale.contiguous(), shift.contiguous(), relu


def pack_bf16(W):
    """(Cout, Cin) f32 -> (Cout, Kp) bf16, K-major, zero padded to a multiple of 64 columns."""
    cout, cin = W.shape
    kp = (cin + _BK - 1) // _BK * _BK
    out = torch.zeros((cout, kp), dtype=torch.bfloat16, device=W.device)
    out[:, :cin] = W.to(torch.bfloat16)
    return out.contiguous(), kp


_GENERATION = 0   # bumped by invalidate_packed(): state changes no tensor version records (CUDA-graph replays)


def invalidate_packed():
    """Drop every cached fold: call after parameters / BatchNorm buffers changed behind autograd's version counters
    (AdaptPointTrainer.step does, for graph replays)."""
    global _GENERATION
    _GENERATION += 1


def _state_key(mods):
    key = [_GENERATION]
    for m in mods:
        for t in list(m.parameters()) + list(m.buffers()):
            key.append((t.data_ptr(), t._version))
    return tuple(key)


def _packed(owner, name, mods, build):
    """Cache folded / packed weights on the module; rebuilt when any parameter or BN buffer changes."""
    cache = owner.__dict__.setdefault("_apb_packed", {})
    key = _state_key(mods)
    hit = cache.get(name)
    if hit is None or hit[0] != key:
        hit = (key, build())
        cache[name] = hit
    return hit[1]


def fusable_reason(grouper, convs, K):
    """None when the fused kernels cover this block, else the reason they do not."""
    from .layers import QueryAndGroup
    if not isinstance(grouper, QueryAndGroup) or not grouper.relative_xyz or grouper.return_only_idx:
        return "grouper is not a relative-xyz ball query"
    if len(convs) not in (1, 2):
        return "more than two conv layers"
    c1 = convs[0][0].out_channels
    if c1 % 8 or c1 > 512 or convs[-1][0].out_channels > 512:
        return "channel count outside the kernel family"
    if len(convs) == 2 and K != 32:
        return "two-layer blocks need nsample == 32"
    return None


@torch.no_grad()
def sa_fused_forward(owner, grouper, convs, query_xyz, support_xyz, features):
    """-> (B, Cout, M) = max_k convs(cat([dp, fj]))  for an eval-mode block (see module docstring)."""
    K = grouper.nsample
    idx = geometry.ball_query(grouper.radius, K, support_xyz, query_xyz)

    def build():
        W1, s1, t1, r1 = fold_conv_block(convs[0])
        w1p = W1[:, :3].t().contiguous()
        if grouper.normalize_dp:
            w1p = (w1p / grouper.radius).contiguous()
        w1f, k1p = pack_bf16(W1[:, 3:])
        pk = dict(w1p=w1p, w1f=w1f, k1p=k1p, s1=s1, t1=t1, r1=r1, c1=W1.shape[0], cin=W1.shape[1] - 3,
                  w2=None, s2=None, t2=None, r2=False, c2=0, k2p=0)
        if len(convs) == 2:
            W2, s2, t2, r2 = fold_conv_block(convs[1])
            w2, k2p = pack_bf16(W2)
            pk.update(w2=w2, k2p=k2p, s2=s2, t2=t2, r2=r2, c2=W2.shape[0])
        return pk

    pk = _packed(owner, "sa", list(convs), build)
    B, N, _ = support_xyz.shape
    M = query_xyz.shape[1]
    feats = features.contiguous().float()
    if feats.shape[1] != pk["cin"]:
        raise ValueError(f"features have {feats.shape[1]} channels, the block expects {pk['cin']}")
    cout = pk["c2"] if pk["w2"] is not None else pk["c1"]
    u_ws = torch.empty((B, N, pk["c1"]), dtype=torch.float32, device=feats.device)
    out = torch.empty((B, cout, M), dtype=torch.float32, device=feats.device)
    _cabi.call("apb_sa_fused_fwd_bf16",
               [support_xyz.contiguous(), query_xyz.contiguous(), feats, idx, pk["w1f"], pk["w1p"], pk["s1"], pk["t1"],
                int(pk["r1"]), pk["w2"], pk["s2"], pk["t2"], int(pk["r2"]), u_ws, out,
                B, N, M, K, pk["cin"], pk["c1"], pk["k1p"], pk["c2"], pk["k2p"]], feats.device)
    return out


@torch.no_grad()
def pwconv_forward(block, x):
    """Eval-mode Conv1d/Conv2d 1x1 (+BN)(+ReLU) block on x (B,Cin,P) or (B,Cin,M,K) -> same rank, Cout channels."""
    def build():
        W, s, t, r = fold_conv_block(block)
        wb, kp = pack_bf16(W)
        return dict(w=wb, kp=kp, s=s, t=t, r=r, cin=W.shape[1], cout=W.shape[0])

    pk = _packed(block, "pw", list(block), build)
    shp = x.shape
    B, cin = shp[0], shp[1]
    P = 1
    for d in shp[2:]:
        P *= d
    xs = x.contiguous().float()
    y = torch.empty((B, pk["cout"]) + tuple(shp[2:]), dtype=torch.float32, device=x.device)
    _cabi.call("apb_pwconv_fwd_bf16", [xs, pk["w"], pk["s"], pk["t"], int(pk["r"]), y, B, cin, pk["cout"], P, pk["kp"]],
               x.device)
    return y
