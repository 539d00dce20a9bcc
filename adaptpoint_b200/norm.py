"""Training-mode BatchNorm (+ReLU) and BatchNorm -> max-over-K on libapb200's streaming kernels (csrc/norm.cu).

The reference builds every conv block as Conv -> BatchNorm{1,2}d -> ReLU (openpoints/models/layers/conv.py:24-103,
norm.py:57-71) and max-pools the neighbour dimension with torch.max (pointnext.py:134); torch runs these as separate
cuDNN / native kernels.  Here
  * `TcBatchNorm1d/2d`  keep nn.BatchNorm{1,2}d's parameters, buffers and state_dict; in training mode on the "tc"
    backend they run two fused passes forward and two backward (optionally with the following ReLU folded in);
  * `bn_max(y, bn, relu)` computes max_k relu?(BN(y)) for y (B,C,M,K) in ONE read pass (batch statistics and the
    per-(b,c,m) max/min/arg together; BN is monotone per channel so max_k BN(y) = BN(max_k y) or BN(min_k y)) and its
    backward in one read + one write pass;
  * `ConvBlock` is the nn.Sequential of a conv block that applies those fusions when it can.
SyncBatchNorm semantics (the reference converts to SyncBN whenever world_size > 1, main.py:27): set
`module.process_group` (see `enable_sync`) and the (C,2) statistics are all-reduced between the two passes.
With the backend at "torch", or in eval mode, everything falls through to the stock torch modules.
"""
import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F
from torch.autograd import Function

from . import _cabi
from . import peer
from . import tcconv


def _dims(x):
    B, C = x.shape[0], x.shape[1]
    P = 1
    for d in x.shape[2:]:
        P *= d
    return B, C, P


def _allreduce(t, group):
    if group is not None and dist.is_initialized() and dist.get_world_size(group) > 1:
        ctx = peer.get(group)
        if ctx is not None and ctx.supports(t):
            ctx.allreduce_(t)            # one single-CTA kernel over NVLink peer memory (csrc/peer.cu)
        else:
            dist.all_reduce(t, group=group)
        return dist.get_world_size(group)
    return 1


class BatchNormActFn(Function):
    """y = relu?(BN_batchstats(x)); x (B,C,*) f32 contiguous."""

    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, momentum, eps, relu, group, stats=None):
        """stats: (C,2) f64 sum / sum of squares of x when the producing GEMM already accumulated them in its epilogue
        (tcconv.conv_with_stats); otherwise one read pass computes them here."""
        xs = x.contiguous()
        B, C, P = _dims(xs)
        dev = xs.device
        if stats is None:
            stats = torch.zeros((C, 2), dtype=torch.float64, device=dev)
            _cabi.call("apb_bn_stats_f32", [xs, stats, B, C, P], dev)
        world = _allreduce(stats, group)
        count = float(B * P * world)
        y = torch.empty_like(xs)
        save_mean = torch.empty(C, dtype=torch.float32, device=dev)
        save_rstd = torch.empty(C, dtype=torch.float32, device=dev)
        _cabi.call("apb_bn_apply_f32", [xs, stats, weight, bias, float(eps), count, int(relu), y, save_mean, save_rstd,
                                        running_mean, running_var, float(momentum), B, C, P], dev)
        ctx.save_for_backward(xs, y if relu else None, weight, save_mean, save_rstd)
        ctx.relu, ctx.group, ctx.count = relu, group, count
        return y

    @staticmethod
    def backward(ctx, gy):
        xs, y, weight, save_mean, save_rstd = ctx.saved_tensors
        B, C, P = _dims(xs)
        dev = xs.device
        g = gy.contiguous().float()
        stats2 = torch.zeros((C, 2), dtype=torch.float64, device=dev)
        _cabi.call("apb_bn_bwd_reduce_f32", [xs, g, y, save_mean, save_rstd, stats2, B, C, P], dev)
        local = stats2.clone() if ctx.group is not None else stats2
        _allreduce(stats2, ctx.group)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(xs)
            _cabi.call("apb_bn_bwd_apply_f32", [xs, g, y, save_mean, save_rstd, weight, stats2, ctx.count, dx, B, C, P],
                       dev)
        # parameter gradients are per-rank sums (the data-parallel all-reduce averages them afterwards)
        dbeta, dgamma = torch.empty_like(save_mean), torch.empty_like(save_mean)
        _cabi.call("apb_sums_split_f32", [local, dbeta, dgamma, C], dev)
        return dx, dgamma, dbeta, None, None, None, None, None, None, None


class BatchNormEvalFn(Function):
    """y = relu?(scale * x + shift) with the running statistics folded into (scale, shift); autograd-capable
    (the AdaptPoint feedback loss back-propagates through the eval-mode classifier, ganloss_cls.py:44-49)."""

    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, eps, relu):
        xs = x.contiguous()
        B, C, P = _dims(xs)
        scale, shift, rstd = (torch.empty(C, dtype=torch.float32, device=xs.device) for _ in range(3))
        _cabi.call("apb_bn_fold_f32", [weight.detach(), bias.detach(), running_mean, running_var, float(eps), scale, shift,
                                       rstd, C], xs.device)
        y = torch.empty_like(xs)
        _cabi.call("apb_affine_f32", [xs, None, scale, shift, int(relu), y, B, C, P], xs.device)
        need_param_grads = weight.requires_grad or bias.requires_grad
        ctx.save_for_backward(xs if need_param_grads else None, y if relu else None, scale, running_mean, rstd)
        ctx.relu = relu
        return y

    @staticmethod
    def backward(ctx, gy):
        xs, y, scale, mean, rstd = ctx.saved_tensors
        g = gy.contiguous().float()
        B, C, P = _dims(g)
        dev = g.device
        dx = dgamma = dbeta = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(g)
            _cabi.call("apb_affine_f32", [g, y, scale, None, 0, dx, B, C, P], dev)
        if (ctx.needs_input_grad[1] or ctx.needs_input_grad[2]) and xs is not None:
            stats2 = torch.zeros((C, 2), dtype=torch.float64, device=dev)
            _cabi.call("apb_bn_bwd_reduce_f32", [xs, g, y, mean, rstd, stats2, B, C, P], dev)
            dbeta, dgamma = torch.empty_like(rstd), torch.empty_like(rstd)
            _cabi.call("apb_sums_split_f32", [stats2, dbeta, dgamma, C], dev)
        return dx, dgamma, dbeta, None, None, None, None


def bnmax_post(vmax, vmin, arg, stats, n, weight, bias, running_mean, running_var, momentum, eps, relu, training):
    """Everything between the per-(b,c,m) max/min and the block's output, in one kernel (apb_bnmax_finalize_f32):
    statistics -> (mean, rstd, scale), running-statistics update, max-or-min by the sign of the scale, affine, ReLU.
    -> out, sel, ksel (B,C,M), scale, mean, rstd (C)."""
    B, C, M = vmax.shape
    dev = vmax.device
    out, sel = torch.empty_like(vmax), torch.empty_like(vmax)
    ksel = torch.empty((B, C, M), dtype=torch.int32, device=dev)
    scale, mean, rstd = (torch.empty(C, dtype=torch.float32, device=dev) for _ in range(3))
    _cabi.call("apb_bnmax_finalize_f32", [stats if training else None, float(n), float(eps), float(momentum or 0.0),
                                          weight.detach(), bias.detach(), running_mean, running_var, vmax, vmin, arg,
                                          int(relu), out, sel, ksel, scale, mean, rstd, B, C, M], dev)
    return out, sel, ksel, scale, mean, rstd


def bnmax_bwd_prep(g, out_relu, sel, scale, mean, rstd, training, group, n):
    """-> gs = g*[out>0]*scale (B,C,M); a, bcoef (C) of the dense BatchNorm-backward terms (None in eval mode); dgamma,
    dbeta (C) per rank.  Two kernels (+ the SyncBN all-reduce of the two channel sums between them)."""
    B, C, M = g.shape
    dev = g.device
    gs = torch.empty_like(g)
    sums = torch.zeros((C, 2), dtype=torch.float64, device=dev)
    _cabi.call("apb_bnmax_bwd_reduce_f32", [g, out_relu, sel, scale, mean, rstd, gs, sums, B, C, M], dev)
    tot = sums
    if training and group is not None and dist.is_initialized() and dist.get_world_size(group) > 1:
        tot = sums.clone()
        _allreduce(tot, group)
    a = bcoef = None
    if training:
        a, bcoef = torch.empty(C, dtype=torch.float32, device=dev), torch.empty(C, dtype=torch.float32, device=dev)
    dgamma, dbeta = torch.empty(C, dtype=torch.float32, device=dev), torch.empty(C, dtype=torch.float32, device=dev)
    _cabi.call("apb_bnmax_bwd_coef_f32", [tot, sums, scale, mean, rstd, float(n), a, bcoef, dgamma, dbeta, C], dev)
    return gs, a, bcoef, dgamma, dbeta


class BNMaxFn(Function):
    """out (B,C,M) = max_k relu?(BN(y))  for y (B,C,M,K); `training` selects batch vs running statistics."""

    @staticmethod
    def forward(ctx, y, weight, bias, running_mean, running_var, momentum, eps, relu, training, group):
        ys = y.contiguous()
        B, C, M, K = ys.shape
        dev = ys.device
        vmax = torch.empty((B, C, M), dtype=torch.float32, device=dev)
        vmin = torch.empty_like(vmax)
        arg = torch.empty((B, C, M), dtype=torch.int32, device=dev)
        stats = torch.zeros((C, 2), dtype=torch.float64, device=dev) if training else None
        _cabi.call("apb_kmax_stats_f32", [ys, vmax, vmin, arg, stats, B, C, M, K], dev)
        n = 1.0
        if training:
            n = float(B * M * K * _allreduce(stats, group))
        out, sel, ksel, scale, mean, rstd = bnmax_post(vmax, vmin, arg, stats, n, weight, bias, running_mean,
                                                       running_var, momentum, eps, relu, training)
        ctx.save_for_backward(ys, sel, ksel, out if relu else None, scale, mean, rstd)
        ctx.cfg = (relu, training, group, n)
        return out

    @staticmethod
    def backward(ctx, g):
        ys, sel, ksel, out, scale, mean, rstd = ctx.saved_tensors
        relu, training, group, n = ctx.cfg
        B, C, M, K = ys.shape
        dev = ys.device
        gs, a, bcoef, dgamma, dbeta = bnmax_bwd_prep(g.contiguous().float(), out, sel, scale, mean, rstd, training, group, n)
        dy = None
        if ctx.needs_input_grad[0]:
            dy = torch.empty_like(ys)
            _cabi.call("apb_bnmax_bwd_f32", [ys if training else None, gs, ksel, a, bcoef, dy, B, C, M, K], dev)
        return dy, dgamma, dbeta, None, None, None, None, None, None, None


def _tc_ok(bn, x):
    return (tcconv.get_conv_backend() == "tc" and x.is_cuda and x.dtype == torch.float32 and bn.affine
            and bn.track_running_stats and bn.momentum is not None)


class _TcBatchNormMixin:
    process_group = None  # set by enable_sync(): SyncBatchNorm semantics over that group

    def forward_act(self, x, relu, stats=None):
        """BatchNorm (+ReLU when `relu`) -- fused kernels in training mode on the tc backend, torch otherwise.
        `stats`: batch statistics of x already accumulated by the producing GEMM (see BatchNormActFn)."""
        if self.training and _tc_ok(self, x):
            y = BatchNormActFn.apply(x, self.weight, self.bias, self.running_mean, self.running_var, self.momentum,
                                     self.eps, relu, self.process_group, stats)
            self.num_batches_tracked.add_(1)
            return y
        if not self.training and _tc_ok(self, x):
            return BatchNormEvalFn.apply(x, self.weight, self.bias, self.running_mean, self.running_var, self.eps, relu)
        y = super().forward(x)
        return F.relu(y, inplace=True) if relu else y

    def forward(self, x):
        return self.forward_act(x, False)


class TcBatchNorm1d(_TcBatchNormMixin, nn.BatchNorm1d):
    pass


class TcBatchNorm2d(_TcBatchNormMixin, nn.BatchNorm2d):
    pass


def bn_max(y, bn, relu):
    """max over the last dim of relu?(bn(y)), y (B,C,M,K); fused when `bn` is a Tc BatchNorm on the tc backend."""
    if isinstance(bn, _TcBatchNormMixin) and _tc_ok(bn, y) and y.dim() == 4:
        out = BNMaxFn.apply(y, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.momentum, bn.eps, relu,
                            bn.training, bn.process_group)
        if bn.training:
            bn.num_batches_tracked.add_(1)
        return out
    z = bn(y)
    if relu:
        z = F.relu(z, inplace=True)
    return torch.max(z, dim=-1, keepdim=False)[0]


class ConvBlock(nn.Sequential):
    """Conv -> [BatchNorm] -> [ReLU] block (same children / state_dict as the reference's nn.Sequential,
    conv.py:24-103); folds the ReLU into the BatchNorm kernels when the norm is a Tc BatchNorm."""

    def forward(self, x):
        mods = list(self)
        i = 0
        stats = None
        while i < len(mods):
            m = mods[i]
            if isinstance(m, _TcBatchNormMixin):
                relu = i + 1 < len(mods) and isinstance(mods[i + 1], nn.ReLU)
                x = m.forward_act(x, relu, stats)
                stats = None
                i += 2 if relu else 1
                continue
            nxt = mods[i + 1] if i + 1 < len(mods) else None
            if (isinstance(m, (tcconv.TcConv1d, tcconv.TcConv2d)) and isinstance(nxt, _TcBatchNormMixin)
                    and nxt.training and _tc_ok(nxt, x)):
                x, stats = tcconv.conv_with_stats(m, x)   # statistics of the conv output from the GEMM epilogue
            else:
                x = m(x)
            i += 1
        return x

    def split_for_pool(self):
        """(conv, bn, relu?) when the block is exactly Conv -> TcBatchNorm [-> ReLU], else None."""
        mods = list(self)
        if len(mods) in (2, 3) and isinstance(mods[1], _TcBatchNormMixin) and \
                (len(mods) == 2 or isinstance(mods[2], nn.ReLU)):
            return mods[0], mods[1], len(mods) == 3
        return None


def enable_sync(module, group=None):
    """SyncBatchNorm semantics for every Tc BatchNorm under `module` (statistics all-reduced over `group`)."""
    group = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
    for m in module.modules():
        if isinstance(m, _TcBatchNormMixin):
            m.process_group = group
    return module
