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
# two-conv blocks (csrc/sa_train2.cu); APB_NO_FUSED_TRAIN2=1 keeps them on the unfused op sequence
ENABLED2 = os.environ.get("APB_NO_FUSED_TRAIN2", "0") != "1"
# Per-step cache of the idx-only backward statistics (cnt, sdp, moments): the InvResMLP blocks of a stage share one
# ball-query result (geometry.ball_query memo), so apb_sa1_bwd_count_f32 runs once per distinct idx per step.
_CNT = {}


def reset_step_cache():
    _CNT.clear()


def _splits(convs):
    out = []
    for blk in convs:
        split = blk.split_for_pool() if isinstance(blk, norm.ConvBlock) else None
        if split is None:
            return None
        out.append(split)
    return out


def usable(mod, grouper, convs, feature_type, reduction, query_p, p, f):
    """True when the block is one or two Conv2d(no bias) -> Tc BatchNorm2d [-> ReLU] layers over a relative-xyz ball query
    with max pooling, on CUDA fp32 tensors the TMA-fed GEMM can address, and the fused path is allowed.  Two-conv blocks
    (csrc/sa_train2.cu) additionally need nsample == 32, a ReLU after the first BatchNorm, C1 <= 256 and C2 <= 512."""
    from .layers import QueryAndGroup
    if not ENABLED or tcconv.get_conv_backend() != "tc" or tcconv.get_conv_precision() != "tf32":
        return False
    if feature_type != "dp_fj" or reduction != "max" or len(convs) not in (1, 2):
        return False
    if len(convs) == 2 and not ENABLED2:
        return False
    if not isinstance(grouper, QueryAndGroup) or not grouper.relative_xyz or grouper.return_only_idx:
        return False
    splits = _splits(convs)
    if splits is None:
        return False
    for conv, bn, _ in splits:
        if conv.bias is not None or not isinstance(conv, tcconv.TcConv2d) or not norm._tc_ok(bn, f):
            return False
        if bn.training != splits[0][1].training or bn.process_group is not splits[0][1].process_group:
            return False
    if f is None or not (f.is_cuda and f.dtype == torch.float32 and p.dtype == torch.float32):
        return False
    N, C1 = p.shape[1], splits[0][0].out_channels
    if N % 4 or C1 % 4 or f.shape[2] != N:
        return False
    if len(splits) == 2:
        C2 = splits[1][0].out_channels
        if grouper.nsample != 32 or not splits[0][2] or C1 > 256 or C2 > 512 or C2 % 4:
            return False
    if splits[0][1].training and torch.is_grad_enabled() and (p.requires_grad or query_p.requires_grad):
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
            cnt, sdp, mom = _idx_moments(ps, qs, idx)
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


def _idx_moments(ps, qs, idx):
    """cnt (B,N), sdp (B,N,3), mom (12) f64 of one ball-query result; cached for the step (see _CNT)."""
    key = (idx.data_ptr(), tuple(idx.shape), ps.data_ptr(), qs.data_ptr())
    hit = _CNT.get(key)
    if hit is None:
        B, N, _ = ps.shape
        M, K = idx.shape[1], idx.shape[2]
        dev = ps.device
        cnt = torch.zeros((B, N), dtype=torch.float32, device=dev)
        sdp = torch.zeros((B, N, 3), dtype=torch.float32, device=dev)
        mom = torch.zeros(12, dtype=torch.float64, device=dev)
        _cabi.call("apb_sa1_bwd_count_f32", [ps, qs, idx, cnt, sdp, mom, B, N, M, K], dev)
        hit = _CNT[key] = (cnt, sdp, mom, idx)     # idx kept alive: its address identifies the entry
    return hit[0], hit[1], hit[2]


class SaTwoLayerFn(Function):
    """max_k act2(BN2(W2 . relu(BN1(W1 . [dp ; f[idx]]))))  (B, C2, M) without any (B, C, M, K) tensor; csrc/sa_train2.cu."""

    @staticmethod
    def forward(ctx, p, q, f, idx, w1, g1, b1, rm1, rv1, mom1, eps1, w2, g2, b2, rm2, rv2, mom2, eps2, relu2, training,
                group, inv_r):
        B, N, _ = p.shape
        M, K = idx.shape[1], idx.shape[2]
        C1, cin = w1.shape[0], w1.shape[1] - 3
        C2 = w2.shape[0]
        dev = f.device
        ps, qs, fs = p.detach().contiguous(), q.detach().contiguous(), f.detach().contiguous().float()
        W1 = w1.detach().reshape(C1, cin + 3).float()
        wp = (W1[:, :3] * inv_r).contiguous()
        wf = W1[:, 3:].contiguous()
        W2 = w2.detach().reshape(C2, C1).float().contiguous()
        if EXACT_GEMM:
            u = torch.einsum("oi,bin->bno", wf, fs).contiguous()
        else:
            wb, kp = tcconv.pack_f32(wf)
            u = torch.empty((B, N, C1), dtype=torch.float32, device=dev)
            _cabi.call("apb_pwconv_ex_tf32", [fs, wb, None, None, 0, u, 1, None, None, B, cin, C1, N, kp], dev)
        s1, t1, mean1, rstd1 = (torch.empty(C1, dtype=torch.float32, device=dev) for _ in range(4))
        n1 = 1.0
        stats1 = stats2 = None
        if training:   # pass 1: batch statistics of the first conv's output
            st = torch.zeros((C1 + C2, 2), dtype=torch.float64, device=dev)     # one fill for both statistics
            stats1, stats2 = st[:C1], st[C1:]
            # closed form from the idx-only moments (cnt, sdp, S1, S2; cached per step, the backward pass needs them too):
            # one pass over the B*N*C1 values of U instead of the B*M*K*C1 gathered ones
            cnt, sdp, mom = _idx_moments(ps, qs, idx)
            _cabi.call("apb_sa_ustats_f32", [u, cnt, sdp, wp, mom, stats1, B, N, C1], dev)
            n1 = float(B * M * K * norm._allreduce(stats1, group))
            _cabi.call("apb_bn_finalize_f32", [stats1, n1, float(eps1), float(mom1 or 0.0), g1.detach(), b1.detach(), rm1,
                                               rv1, s1, t1, mean1, rstd1, C1], dev)
        else:
            _cabi.call("apb_bn_fold_f32", [g1.detach(), b1.detach(), rm1, rv1, float(eps1), s1, t1, rstd1, C1], dev)
            mean1 = rm1.detach().float().clone()
        w2b, k2p = tcconv.pack_f32(W2)
        vmax = torch.empty((B, C2, M), dtype=torch.float32, device=dev)
        vmin = torch.empty_like(vmax)
        arg = torch.empty((B, C2, M), dtype=torch.int32, device=dev)
        _cabi.call("apb_sa2_fwd_tf32", [u, ps, qs, idx, wp, s1, t1, 1, w2b, vmax, vmin, arg, stats2, B, N, M, K, C1, C2,
                                        k2p], dev)
        n2 = 1.0
        if training:
            n2 = float(B * M * K * norm._allreduce(stats2, group))
        out, sel, ksel, scale2, mean2, rstd2 = norm.bnmax_post(vmax, vmin, arg, stats2, n2, g2, b2, rm2, rv2, mom2, eps2,
                                                               relu2, training)
        ksel_pm = torch.empty((B, M, C2), dtype=torch.uint8, device=dev).copy_(ksel.transpose(1, 2))   # one kernel
        ctx.save_for_backward(ps, qs, fs, idx, u, sel, ksel_pm, out if relu2 else None, scale2, mean2, rstd2, wp, wf, W2, s1,
                              t1, mean1, rstd1)
        ctx.cfg = (relu2, training, group, n1, n2, inv_r, tuple(w1.shape), tuple(w2.shape))
        return out

    @staticmethod
    def backward(ctx, g):
        (ps, qs, fs, idx, u, sel, ksel_pm, out, scale2, mean2, rstd2, wp, wf, W2, s1, t1, mean1, rstd1) = ctx.saved_tensors
        relu2, training, group, n1, n2, inv_r, w1shape, w2shape = ctx.cfg
        B, N, _ = ps.shape
        M, K = idx.shape[1], idx.shape[2]
        C1, cin = wf.shape
        C2 = W2.shape[0]
        dev = fs.device
        gs, a2, b2c, dgamma2, dbeta2 = norm.bnmax_bwd_prep(g.contiguous().float(), out, sel, scale2, mean2, rstd2, training,
                                                           group, n2)
        gs_pm = gs.transpose(1, 2).contiguous()                          # (B, M, C2)
        need_w = ctx.needs_input_grad[4] or ctx.needs_input_grad[11]
        need_bn1 = ctx.needs_input_grad[5] or ctx.needs_input_grad[6]
        kc2, c1pad = (C2 + 31) // 32, (C1 + 31) // 32 * 32
        want_sums = training or need_w or need_bn1
        # one fp32 and one fp64 zero fill for everything the kernels accumulate into
        kw = kc2 * 32 + (c1pad if training else 0)
        n_du, n_wb, n_dm = B * N * C1, C1 * kw, ((C2 + (C1 if training else 0)) * c1pad if need_w else 0)
        z32 = torch.zeros(n_du + n_wb + n_dm, dtype=torch.float32, device=dev)
        du_pm = z32[:n_du].view(B, N, C1)
        wb = z32[n_du:n_du + n_wb].view(C1, kw)
        dmat = z32[n_du + n_wb:].view(-1, c1pad) if need_w else None
        sums1 = sdzdp = sh1 = usdp = None
        if want_sums:
            z64 = torch.zeros(9 * C1, dtype=torch.float64, device=dev)
            sums1, sdzdp = z64[:2 * C1].view(C1, 2), z64[2 * C1:5 * C1].view(C1, 3)
            sh1, usdp = z64[5 * C1:6 * C1], z64[6 * C1:].view(C1, 3)
        wb[:, :C2] = W2.t()
        vvec = bw = None
        if training:   # dense BatchNorm-2 terms: dh1 += W2^T a2 + (W2^T diag(b2) W2) h1
            bw = b2c.view(C2, 1) * W2
            wb[:, kc2 * 32:kc2 * 32 + C1] = W2.t() @ bw
            vvec = (W2.t() @ a2).contiguous()
        dxyz = torch.zeros_like(ps) if ctx.needs_input_grad[0] else None
        dq = torch.zeros_like(qs) if ctx.needs_input_grad[1] else None
        _cabi.call("apb_sa2_bwd_rows_tf32", [u, ps, qs, idx, wp, s1, t1, 1, gs_pm, ksel_pm, wb, wb.shape[1], kc2, vvec, mean1,
                                             rstd1, du_pm, sums1, sdzdp, sh1, dxyz, dq, B, N, M, K, C1, C2], dev)
        du = du_pm.transpose(1, 2).contiguous()                          # (B, C1, N) for the dgrad / wgrad GEMMs
        dgamma1 = dbeta1 = None
        dwp = None
        if training:
            tot1 = sums1
            if group is not None and torch.distributed.is_initialized() and torch.distributed.get_world_size(group) > 1:
                tot1 = sums1.clone()
                norm._allreduce(tot1, group)
            a1, b1c = (torch.empty(C1, dtype=torch.float32, device=dev) for _ in range(2))
            dgamma1, dbeta1 = (torch.empty(C1, dtype=torch.float32, device=dev) for _ in range(2))
            _cabi.call("apb_bnmax_bwd_coef_f32", [tot1, sums1, s1, mean1, rstd1, float(n1), a1, b1c, dgamma1, dbeta1, C1], dev)
            cnt, sdp, mom = _idx_moments(ps, qs, idx)
            _cabi.call("apb_sa1_bwd_dense_f32", [u, cnt, sdp, a1, b1c, wp, du, usdp, B, N, C1], dev)
            if need_w:   # dWp = s1 * sum dz1 dp + a1 S1^T + b1 * (U^T sdp + Wp S2)
                cd = torch.stack([s1, a1, b1c, wp[:, 0], wp[:, 1], wp[:, 2]]).double()      # one conversion for all
                dwp = torch.addmm(usdp, cd[3:].t(), mom[3:].view(3, 3)).mul_(cd[2].view(C1, 1))
                dwp.addcmul_(cd[0].view(C1, 1), sdzdp).add_(torch.outer(cd[1], mom[:3]))
        elif need_w:
            dwp = s1.double().view(C1, 1) * sdzdp                          # sum_r (s1 * dz1) dp
        if not training and need_bn1:
            dbeta1, dgamma1 = (torch.empty(C1, dtype=torch.float32, device=dev) for _ in range(2))
            _cabi.call("apb_sums_split_f32", [sums1, dbeta1, dgamma1, C1], dev)
        df = dw1 = dw2 = None
        if ctx.needs_input_grad[2]:
            df = (torch.einsum("oi,bon->bin", wf, du) if EXACT_GEMM
                  else tcconv._gemm_rows(du, wf.t().contiguous(), None, B, C1, cin, N))
        if need_w:
            dwf = torch.einsum("bon,bin->oi", du, fs) if EXACT_GEMM else tcconv.wgrad(du, fs, B, cin, C1, N)
            dw1 = torch.cat([(dwp * inv_r).float(), dwf], dim=1).view(w1shape)
            _cabi.call("apb_sa2_wgrad_tf32", [u, ps, qs, idx, wp, s1, t1, 1, gs_pm, ksel_pm, dmat, int(training), B, N, M, K,
                                              C1, C2], dev)
            dw2 = dmat[:C2, :C1]
            if training:   # dense terms of dy2: diag(b2) W2 (sum h1 h1^T) + a2 (sum h1)^T
                dw2 = torch.addmm(dw2, bw, dmat[C2:C2 + C1, :C1]).addr_(a2, sh1.float())
            dw2 = dw2.reshape(w2shape)
        # BatchNorm parameter gradients are per-rank sums (the data-parallel all-reduce averages them afterwards)
        return (dxyz, dq, df, None, dw1, dgamma1, dbeta1, None, None, None, None, dw2, dgamma2, dbeta2, None, None, None,
                None, None, None, None, None)


def sa_two_layer(mod, grouper, convs, query_p, p, f):
    """-> (B, C2, M) for a two-conv block, see `usable`."""
    (conv1, bn1, _), (conv2, bn2, relu2) = _splits(convs)
    idx = geometry.ball_query(grouper.radius, grouper.nsample, p, query_p)
    inv_r = 1.0 / grouper.radius if grouper.normalize_dp else 1.0
    out = SaTwoLayerFn.apply(p, query_p, f, idx, conv1.weight, bn1.weight, bn1.bias, bn1.running_mean, bn1.running_var,
                             bn1.momentum, bn1.eps, conv2.weight, bn2.weight, bn2.bias, bn2.running_mean, bn2.running_var,
                             bn2.momentum, bn2.eps, relu2, bn1.training, bn1.process_group, inv_r)
    if bn1.training:
        bn1.num_batches_tracked.add_(1)
        bn2.num_batches_tracked.add_(1)
    return out


def sa_one_layer(mod, grouper, convs, query_p, p, f):
    """-> (B, C1, M) = max_k act(BN(conv(cat([dp, fj]))))  for a one-conv block, see `usable`."""
    if len(convs) == 2:
        return sa_two_layer(mod, grouper, convs, query_p, p, f)
    conv, bn, relu = convs[0].split_for_pool()
    idx = geometry.ball_query(grouper.radius, grouper.nsample, p, query_p)
    inv_r = 1.0 / grouper.radius if grouper.normalize_dp else 1.0
    out = SaOneLayerFn.apply(p, query_p, f, idx, conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                             bn.momentum, bn.eps, relu, bn.training, bn.process_group, inv_r)
    if bn.training:
        bn.num_batches_tracked.add_(1)
    return out
