"""TEST INFRASTRUCTURE: torch (float64) restatements of the C-ABI entry points behind adaptpoint_b200.fused_train's
two-conv set abstraction (csrc/sa_train2.cu and the norm.cu / sa_train.cu helpers it shares with the one-conv path).

Two uses:
  * tests/test_fused_train_cpu.py swaps `_cabi.call` for `call` below and checks the host-side algebra of
    SaTwoLayerFn (closed-form BatchNorm backward, Gram-matrix weight gradient, ...) against plain torch autograd of the
    reference op chain (pointnext.py:164-166) -- no GPU needed;
  * tests/test_fused_train_gpu.py runs every sa_train2.cu kernel and this restatement on the same inputs.
Each function takes the same argument list as the C entry point (minus the stream) and writes its outputs in place.
"""
import torch


def _rows(u, xyz, new_xyz, idx, wp):
    """y1 (B,M,K,C1) f64 and dp (B,M,K,3) f64 of every (query, neighbour) row."""
    B = u.shape[0]
    j = idx.long()
    bi = torch.arange(B, device=u.device).view(B, 1, 1)
    dp = xyz.double()[bi, j] - new_xyz.double()[:, :, None, :]
    y1 = u.double()[bi, j] + dp @ wp.double().t()
    return y1, dp


def _minmax(y, vmax, vmin, arg):
    """y (B,M,K,C) -> vmax/vmin (B,C,M), arg = kmax | kmin << 16 (first k wins ties)."""
    K = y.shape[2]
    mx, mn = y.max(dim=2).values, y.min(dim=2).values
    ks = torch.arange(K, device=y.device).view(1, 1, K, 1)
    kx = torch.where(y == mx[:, :, None, :], ks, K).min(dim=2).values
    kn = torch.where(y == mn[:, :, None, :], ks, K).min(dim=2).values
    vmax.copy_(mx.permute(0, 2, 1))
    vmin.copy_(mn.permute(0, 2, 1))
    arg.copy_((kx | (kn << 16)).permute(0, 2, 1).to(arg.dtype))


def apb_group_ystats_f32(u, xyz, new_xyz, idx, wp, vmax, vmin, arg, stats, B, N, M, K, C):
    y, _ = _rows(u, xyz, new_xyz, idx, wp)
    if vmax is not None:
        _minmax(y, vmax, vmin, arg)
    if stats is not None:
        stats[:, 0] += y.sum(dim=(0, 1, 2))
        stats[:, 1] += (y * y).sum(dim=(0, 1, 2))


def apb_bn_finalize_f32(stats, count, eps, momentum, gamma, beta, rm, rv, scale, shift, mean, rstd, C):
    m = stats[:, 0] / count
    var = (stats[:, 1] / count - m * m).clamp_min(0.0)
    r = 1.0 / torch.sqrt(var + eps)
    if rm is not None:
        rm.copy_((1 - momentum) * rm + momentum * m.float())
        rv.copy_((1 - momentum) * rv + momentum * (var * (count / (count - 1) if count > 1 else 1.0)).float())
    mean.copy_(m)
    rstd.copy_(r)
    scale.copy_(gamma.double() * r)
    shift.copy_(beta.double() - m * gamma.double() * r)


def apb_bn_fold_f32(gamma, beta, rm, rv, eps, scale, shift, rstd, C):
    r = 1.0 / torch.sqrt(rv.double() + eps)
    rstd.copy_(r)
    scale.copy_(gamma.double() * r)
    shift.copy_(beta.double() - rm.double() * gamma.double() * r)


def _h1(u, xyz, new_xyz, idx, wp, s1, t1, relu1):
    y1, dp = _rows(u, xyz, new_xyz, idx, wp)
    z1 = y1 * s1.double() + t1.double()
    return y1, dp, z1, (z1.clamp_min(0.0) if relu1 else z1)


def apb_sa2_fwd_tf32(u, xyz, new_xyz, idx, wp, s1, t1, relu1, w2, vmax, vmin, arg, stats, B, N, M, K, C1, C2, K2p):
    _, _, _, h1 = _h1(u, xyz, new_xyz, idx, wp, s1, t1, relu1)
    y2 = h1 @ w2.double()[:, :C1].t()
    _minmax(y2, vmax, vmin, arg)
    if stats is not None:
        stats[:, 0] += y2.sum(dim=(0, 1, 2))
        stats[:, 1] += (y2 * y2).sum(dim=(0, 1, 2))


def apb_bnmax_finalize_f32(stats, n, eps, momentum, weight, bias, rm, rv, vmax, vmin, arg, relu, out, sel, ksel, scale,
                           mean, rstd, B, C, M):
    if stats is not None:
        m = stats[:, 0] / n
        var = (stats[:, 1] / n - m * m).clamp_min(0.0)
        r = 1.0 / torch.sqrt(var + eps)
        if rm is not None:
            rm.copy_((1 - momentum) * rm + momentum * m.float())
            rv.copy_((1 - momentum) * rv + momentum * (var * (n / (n - 1) if n > 1 else 1.0)).float())
    else:
        m, r = rm.double(), 1.0 / torch.sqrt(rv.double() + eps)
    sc = weight.double() * r
    sh = bias.double() - m * sc
    pos = (sc >= 0).view(1, C, 1)
    v = torch.where(pos, vmax.double(), vmin.double())
    o = v * sc.view(1, C, 1) + sh.view(1, C, 1)
    out.copy_(o.clamp_min(0.0) if relu else o)
    sel.copy_(v)
    ksel.copy_(torch.where(pos, arg & 0xFFFF, arg >> 16))
    scale.copy_(sc)
    mean.copy_(m)
    rstd.copy_(r)


def apb_bnmax_bwd_reduce_f32(g, out_relu, sel, scale, mean, rstd, gs, sums, B, C, M):
    gv = g.double()
    if out_relu is not None:
        gv = torch.where(out_relu > 0, gv, torch.zeros_like(gv))
    gs.copy_(gv * scale.double().view(1, C, 1))
    xhat = (sel.double() - mean.double().view(1, C, 1)) * rstd.double().view(1, C, 1)
    sums[:, 0] += gv.sum(dim=(0, 2))
    sums[:, 1] += (gv * xhat).sum(dim=(0, 2))


def apb_bnmax_bwd_coef_f32(tot, local, scale, mean, rstd, count, a, bcoef, dgamma, dbeta, C):
    dbeta.copy_(local[:, 0])
    dgamma.copy_(local[:, 1])
    if a is not None:
        bc = -scale.double() * rstd.double() * tot[:, 1] / count
        bcoef.copy_(bc)
        a.copy_(-scale.double() * tot[:, 0] / count - bc * mean.double())


def _onehot(gs_pm, ksel_pm, B, M, K, C2):
    """(B,M,K,C2) f64: gs at k == ksel, else 0."""
    ks = torch.arange(K, device=gs_pm.device).view(1, 1, K, 1)
    hit = ksel_pm.view(B, M, 1, C2).long() == ks
    return torch.where(hit, gs_pm.double().view(B, M, 1, C2), torch.zeros((), dtype=torch.float64, device=gs_pm.device))


def apb_sa2_bwd_rows_tf32(u, xyz, new_xyz, idx, wp, s1, t1, relu1, gs_pm, ksel_pm, wb, Kp, kc_sparse, vvec, mean1, rstd1,
                          du_pm, sums1, sdzdp, sh1, dxyz, dq, B, N, M, K, C1, C2):
    y1, dp, z1, h1 = _h1(u, xyz, new_xyz, idx, wp, s1, t1, relu1)
    oh = _onehot(gs_pm, ksel_pm, B, M, K, C2)
    wbd = wb.double()
    dh1 = oh @ wbd[:, :C2].t()
    if Kp > kc_sparse * 32:
        dh1 = dh1 + h1 @ wbd[:, kc_sparse * 32:kc_sparse * 32 + C1].t()
    if vvec is not None:
        dh1 = dh1 + vvec.double()
    dz1 = torch.where(z1 > 0, dh1, torch.zeros_like(dh1)) if relu1 else dh1
    ds = dz1 * s1.double()
    j = idx.long().view(B, M * K)
    acc = torch.zeros((B, N, C1), dtype=torch.float64, device=u.device)
    acc.scatter_add_(1, j.unsqueeze(-1).expand(-1, -1, C1), ds.view(B, M * K, C1))
    du_pm += acc.to(du_pm.dtype)
    if sums1 is not None:
        xhat = (y1 - mean1.double()) * rstd1.double()
        sums1[:, 0] += dz1.sum(dim=(0, 1, 2))
        sums1[:, 1] += (dz1 * xhat).sum(dim=(0, 1, 2))
        if sdzdp is not None:
            sdzdp += torch.einsum("bmkc,bmkj->cj", dz1, dp)
        if sh1 is not None:
            sh1 += h1.sum(dim=(0, 1, 2))
    if dxyz is not None or dq is not None:
        gdp = ds @ wp.double()      # (B,M,K,3)
        if dxyz is not None:
            acc3 = torch.zeros((B, N, 3), dtype=torch.float64, device=u.device)
            acc3.scatter_add_(1, j.unsqueeze(-1).expand(-1, -1, 3), gdp.view(B, M * K, 3))
            dxyz += acc3.to(dxyz.dtype)
        if dq is not None:
            dq -= gdp.sum(dim=2).to(dq.dtype)


def apb_sa1_bwd_count_f32(xyz, new_xyz, idx, cnt, sdp, mom, B, N, M, K):
    j = idx.long().view(B, M * K)
    bi = torch.arange(B, device=xyz.device).view(B, 1, 1)
    dp = (xyz.double()[bi, idx.long()] - new_xyz.double()[:, :, None, :]).view(B, M * K, 3)
    c = torch.zeros((B, N), dtype=torch.float64, device=xyz.device)
    c.scatter_add_(1, j, torch.ones_like(j, dtype=torch.float64))
    s = torch.zeros((B, N, 3), dtype=torch.float64, device=xyz.device)
    s.scatter_add_(1, j.unsqueeze(-1).expand(-1, -1, 3), dp)
    cnt += c.to(cnt.dtype)
    sdp += s.to(sdp.dtype)
    mom[:3] += dp.sum(dim=(0, 1))
    mom[3:] += torch.einsum("bri,brj->ij", dp, dp).reshape(9)


def apb_sa1_bwd_dense_f32(u, cnt, sdp, a, bc, wp, du_cm, usdp, B, N, C):
    ud, cn, sd = u.double(), cnt.double().unsqueeze(-1), sdp.double()
    d = cn * a.double() + bc.double() * (cn * ud + sd @ wp.double().t())     # (B,N,C)
    du_cm += d.permute(0, 2, 1).to(du_cm.dtype)
    usdp += torch.einsum("bnc,bnj->cj", ud, sd)


def apb_sa_ustats_f32(u, cnt, sdp, wp, mom, stats, B, N, C):
    ud, cn, sd, w = u.double(), cnt.double().unsqueeze(-1), sdp.double(), wp.double()
    ws = sd @ w.t()                                            # (B,N,C): Wp[c,:] . sdp[n]
    stats[:, 0] += (cn * ud).sum(dim=(0, 1)) + w @ mom[:3]
    stats[:, 1] += (cn * ud * ud + 2 * ud * ws).sum(dim=(0, 1)) + ((w @ mom[3:].view(3, 3)) * w).sum(dim=1)


def apb_sums_split_f32(sums, first, second, C):
    first.copy_(sums[:, 0])
    second.copy_(sums[:, 1])


def apb_sa2_wgrad_tf32(u, xyz, new_xyz, idx, wp, s1, t1, relu1, gs_pm, ksel_pm, d_out, with_gram, B, N, M, K, C1, C2):
    _, _, _, h1 = _h1(u, xyz, new_xyz, idx, wp, s1, t1, relu1)
    oh = _onehot(gs_pm, ksel_pm, B, M, K, C2)
    d_out[:C2, :C1] += torch.einsum("bmkd,bmkc->dc", oh, h1).to(d_out.dtype)
    if with_gram:
        d_out[C2:C2 + C1, :C1] += torch.einsum("bmkd,bmkc->dc", h1, h1).to(d_out.dtype)


def call(name, args, device):
    """Drop-in for adaptpoint_b200._cabi.call."""
    fn = globals().get(name)
    if fn is None:
        raise KeyError(f"no restatement of {name}")
    fn(*args)


def reference_block(p, q, f, idx, w1, g1, b1, w2, g2, b2, relu2, training, inv_r, rm1=None, rv1=None, rm2=None, rv2=None,
                    eps=1e-5):
    """The reference's op chain with stock torch ops (group -> cat -> conv -> BN -> ReLU -> conv -> BN [-> ReLU] -> max),
    autograd-capable; all tensors float64."""
    import torch.nn.functional as F
    B = p.shape[0]
    j = idx.long()
    bi = torch.arange(B, device=p.device).view(B, 1, 1)
    dp = (p[bi, j] - q[:, :, None, :]) * inv_r                  # (B,M,K,3)
    fj = f.permute(0, 2, 1)[bi, j]                              # (B,M,K,Cin)
    x = torch.cat([dp, fj], dim=-1).permute(0, 3, 1, 2)         # (B,3+Cin,M,K)
    y = F.conv2d(x, w1)
    y = F.batch_norm(y, rm1, rv1, g1, b1, training, 0.1, eps)
    y = F.relu(y)
    y = F.conv2d(y, w2)
    y = F.batch_norm(y, rm2, rv2, g2, b2, training, 0.1, eps)
    if relu2:
        y = F.relu(y)
    return y.max(dim=-1).values
