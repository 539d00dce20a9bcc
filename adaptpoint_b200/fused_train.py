"""Set abstraction of ONE-conv blocks under autograd without the (B,C,M,K) tensors (csrc/sa_train.cu).

Stands in, in TRAINING mode (BatchNorm batch statistics, SyncBN-capable) and in eval mode under autograd, for the op chain
  SetAbstraction.forward      pointnext.py:164-166   grouper -> cat([dp, fj]) -> convs -> max over K      (sa_layers = 1)
  LocalAggregation.forward    pointnext.py:66-77     same, query set == support set
  with convs = Conv2d 1x1 (no bias) -> BatchNorm2d [-> ReLU]                                conv.py:24-60
i.e. every aggregation block of PointNeXt-B / L / XL (BASELINE.json configs[3], [4]).  The algebra is in the header of
csrc/sa_train.cu: U = Wf . f per support point (TMA-fed TF32 GEMM on B*N rows), one gather pass for max/min over K
plus the batch statistics, and a backward pass whose dense BatchNorm terms collapse to per-support-point closed forms,
so dgrad / wgrad run on B*N rows instead of B*M*K.  Numerics: TF32 operands in the U GEMM and its two gradients (the
training-path precision of tcconv.py), everything else fp32 / fp64 sums.
"""
import os

import torch
from torch.autograd import Function

from . import _cabi, geometry, norm, tcconv

# benchmarks / tests flip this (or set APB_NO_FUSED_TRAIN=1) to time the unfused op sequence on the same kernels
ENABLED = os.environ.get("APB_NO_FUSED_TRAIN", "0") != "1"
# TEST-ONLY: evaluate the three dense contractions (U = Wf.f, df = Wf^T.dU, dWf = dU.f^T) with fp32 einsums instead of
# the TF32 tensor-core kernels, so that the gather / statistics / closed-form backward kernels of csrc/sa_train.cu can be
# compared element by element with the unfused fp32 op sequence (TF32 noise flips a fraction of the max-over-K routes,
# which moves whole gradient entries between support points).
EXACT_GEMM = False
# Per-step cache of the idx-only backward statistics (cnt, sdp, moments): the InvResMLP blocks of a stage share one
# ball-query result (geometry.ball_query memo), so apb_sa1_bwd_count_f32 runs once per distinct idx per step.
_CNT = {}


def reset_step_cache():
    _CNT.clear()


def usable(mod, grouper, convs, feature_type, reduction, query_p, p, f):
    """True when the block is one Conv2d(no bias) -> Tc BatchNorm2d [-> ReLU] over a relative-xyz ball query with max
    pooling, on CUDA fp32 tensors the TMA-fed GEMM can address, and the fused path is allowed."""
    from .layers import QueryAndGroup
    if not ENABLED or tcconv.get_conv_backend() != "tc" or tcconv.get_conv_precision() != "tf32":
        return False
    if feature_type != "dp_fj" or reduction != "max" or len(convs) != 1:
        return False
    if not isinstance(grouper, QueryAndGroup) or not grouper.relative_xyz or grouper.return_only_idx:
        return False
    split = convs[0].split_for_pool() if isinstance(convs[0], norm.ConvBlock) else None
    if split is None:
        return False
    conv, bn, _ = split
    if conv.bias is not None or not isinstance(conv, tcconv.TcConv2d) or not norm._tc_ok(bn, f):
        return False
    if f is None or not (f.is_cuda and f.dtype == torch.float32 and p.dtype == torch.float32):
        return False
    N, C1 = p.shape[1], conv.out_channels
    if N % 4 or C1 % 4 or f.shape[2] != N:
        return False
    if bn.training and torch.is_grad_enabled() and (p.requires_grad or query_p.requires_grad):
        return False      # coordinate gradients through batch statistics: not on the fused path (no model needs them)
    return True


class SaOneLayerFn(Function):
    @staticmethod
    def forward(ctx, p, q, f, idx, conv_w, gamma, beta, running_mean, running_var, momentum, eps, relu, training, group,
                inv_r):
        B, N, _ = p.shape
        M, K = idx.shape[1], idx.shape[2]
        C1, cin = conv_w.shape[0], conv_w.shape[1] - 3
        dev = f.device
        ps, qs, fs = p.detach().contiguous(), q.detach().contiguous(), f.detach().contiguous().float()
        W = conv_w.detach().reshape(C1, cin + 3).float()
        wp = (W[:, :3] * inv_r).contiguous()
        wf = W[:, 3:].contiguous()
        wb, kp = tcconv.pack_f32(wf)
        if EXACT_GEMM:
            u = torch.einsum("oi,bin->bno", wf, fs).contiguous()
        else:
            u = torch.empty((B, N, C1), dtype=torch.float32, device=dev)
            _cabi.call("apb_pwconv_ex_tf32", [fs, wb, None, None, 0, u, 1, None, None, B, cin, C1, N, kp], dev)
        vmax = torch.empty((B, C1, M), dtype=torch.float32, device=dev)
        vmin = torch.empty_like(vmax)
        arg = torch.empty((B, C1, M), dtype=torch.int32, device=dev)
        stats = torch.zeros((C1, 2), dtype=torch.float64, device=dev) if training else None
        _cabi.call("apb_group_ystats_f32", [u, ps, qs, idx, wp, vmax, vmin, arg, stats, B, N, M, K, C1], dev)
        n = 1.0
        if training:
            n = float(B * M * K * norm._allreduce(stats, group))
        out, sel, ksel, scale, mean, rstd = norm.bnmax_post(vmax, vmin, arg, stats, n, gamma, beta, running_mean,
                                                            running_var, momentum, eps, relu, training)
        ctx.save_for_backward(ps, qs, fs, idx, u, sel, ksel, out if relu else None, scale, mean, rstd, wp, wf)
        ctx.cfg = (relu, training, group, n, inv_r, tuple(conv_w.shape))
        return out

    @staticmethod
    def backward(ctx, g):
        ps, qs, fs, idx, u, sel, ksel, out, scale, mean, rstd, wp, wf = ctx.saved_tensors
        relu, training, group, n, inv_r, wshape = ctx.cfg
        B, N, _ = ps.shape
        M, K = idx.shape[1], idx.shape[2]
        C1, cin = wf.shape
        dev = fs.device
        gs, a, bcoef, sgx, sg = norm.bnmax_bwd_prep(g.contiguous().float(), out, sel, scale, mean, rstd, training, group, n)
        need_w = ctx.needs_input_grad[4]
        du = torch.zeros((B, C1, N), dtype=torch.float32, device=dev)
        dwp = torch.zeros((C1, 3), dtype=torch.float64, device=dev) if need_w else None
        dxyz = torch.zeros_like(ps) if ctx.needs_input_grad[0] else None
        dq = torch.zeros_like(qs) if ctx.needs_input_grad[1] else None
        _cabi.call("apb_sa1_bwd_sparse_f32", [gs, ksel, idx, ps, qs, wp, du, dwp, dxyz, dq, B, N, M, K, C1], dev)
        if training:
            key = (idx.data_ptr(), tuple(idx.shape), ps.data_ptr(), qs.data_ptr())
            hit = _CNT.get(key)
            if hit is None:
                cnt = torch.zeros((B, N), dtype=torch.float32, device=dev)
                sdp = torch.zeros((B, N, 3), dtype=torch.float32, device=dev)
                mom = torch.zeros(12, dtype=torch.float64, device=dev)
                _cabi.call("apb_sa1_bwd_count_f32", [ps, qs, idx, cnt, sdp, mom, B, N, M, K], dev)
                _CNT[key] = (cnt, sdp, mom, idx)     # idx kept alive: its address identifies the entry
            else:
                cnt, sdp, mom, _ = hit
            usdp = torch.zeros((C1, 3), dtype=torch.float64, device=dev)
            _cabi.call("apb_sa1_bwd_dense_f32", [u, cnt, sdp, a, bcoef, wp, du, usdp, B, N, C1], dev)
            if need_w:
                s1, s2 = mom[:3], mom[3:].view(3, 3)
                dwp = dwp + a.double().view(C1, 1) * s1.view(1, 3) + bcoef.double().view(C1, 1) * (usdp + wp.double() @ s2)
        df = dw = None
        if ctx.needs_input_grad[2]:
            df = (torch.einsum("oi,bon->bin", wf, du) if EXACT_GEMM
                  else tcconv._gemm_rows(du, wf.t().contiguous(), None, B, C1, cin, N))
        if need_w:
            if EXACT_GEMM:
                dwf = torch.einsum("bon,bin->oi", du, fs)
            else:
                dwf = tcconv.wgrad(du, fs, B, cin, C1, N)
            dw = torch.cat([(dwp * inv_r).float(), dwf], dim=1).view(wshape)
        # BatchNorm parameter gradients are per-rank sums (the data-parallel all-reduce averages them afterwards)
        return dxyz, dq, df, None, dw, sgx, sg, None, None, None, None, None, None, None, None


def sa_one_layer(mod, grouper, convs, query_p, p, f):
    """-> (B, C1, M) = max_k act(BN(conv(cat([dp, fj]))))  for a one-conv block, see `usable`."""
    conv, bn, relu = convs[0].split_for_pool()
    idx = geometry.ball_query(grouper.radius, grouper.nsample, p, query_p)
    inv_r = 1.0 / grouper.radius if grouper.normalize_dp else 1.0
    out = SaOneLayerFn.apply(p, query_p, f, idx, conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                             bn.momentum, bn.eps, relu, bn.training, bn.process_group, inv_r)
    if bn.training:
        bn.num_batches_tracked.add_(1)
    return out
