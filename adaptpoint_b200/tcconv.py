"""1x1 convolutions (the "shared MLP" layers) on the tensor cores, forward AND backward, for training.

nn.Conv1d / nn.Conv2d with kernel 1 is what every conv of the path is in the reference
(openpoints/models/layers/conv.py:8-21: Conv1d/Conv2d wrappers; create_convblock1d/2d :24-103); torch runs them
through cuDNN.  `TcConv1d` / `TcConv2d` keep the parameters, names and state_dict of nn.Conv{1,2}d and, when the
backend switch is "tc", run
    y  = W x (+b)            apb_pwconv_fwd_tf32 | _bf16   (tf32 or bf16 operands, fp32 accumulate, tcgen05)
    gx = W^T gy              the same kernel on the transposed weights
    gW = gy x^T              apb_pwconv_wgrad_tf32 (TF32, both operands TMA-fed)
Numerics: tf32 (default) ~1e-3 relative per layer; bf16 2e-2 (BASELINE.json north_star, "bf16 MLP path").  "tc" is
the product path and the default for CUDA tensors; the switch at "torch" turns the layers back into plain
nn.Conv{1,2}d -- a TEST-ONLY strict-fp32 mode (tests/, the golden parity runs); CPU tensors (the op-patched CPU
tests and the CPU baseline of bench.py) always take the nn.Conv path, there is no CPU product path.
"""
import contextlib

import torch
import torch.nn as nn
from torch.autograd import Function

from . import _cabi
from .fused import pack_bf16

_BACKEND = "tc"
_PRECISION = "tf32"   # operands of y = W x and gx = W^T gy on the "tc" backend: "tf32" (default) or "bf16"
_MAX_N = 512  # accumulator columns of one kernel call


def set_conv_precision(name):
    """'tf32' (10-bit mantissa operands: what the reference's cuDNN convs use by default on Ampere+) or 'bf16'
    (BASELINE.json's "bf16 MLP path", 2e-2).  bf16 operand rounding flips enough max-pool / ReLU routing decisions
    in PointNeXt to turn the weight gradients of the early stages ~40 degrees (cos 0.75 vs fp32, measured against
    an exact torch emulation in round 1), tf32 keeps cos > 0.999, at the same HBM-bound speed."""
    global _PRECISION
    if name not in ("tf32", "bf16"):
        raise ValueError(name)
    _PRECISION = name


def get_conv_precision():
    return _PRECISION


def set_conv_backend(name):
    """'tc' (libapb200 tensor-core kernels: the product path, default) or 'torch' (cuDNN fp32: test-only strict mode)."""
    global _BACKEND
    if name not in ("torch", "tc"):
        raise ValueError(name)
    _BACKEND = name


def get_conv_backend():
    return _BACKEND


@contextlib.contextmanager
def conv_backend(name):
    prev = _BACKEND
    set_conv_backend(name)
    try:
        yield
    finally:
        set_conv_backend(prev)


def pack_f32(W):
    """(Cout, Cin) f32 -> (Cout, Kp) f32, K-major, zero padded to a multiple of 32 columns."""
    cout, cin = W.shape
    kp = (cin + 31) // 32 * 32
    if kp == cin:
        return W.contiguous(), kp
    out = torch.zeros((cout, kp), dtype=torch.float32, device=W.device)
    out[:, :cin] = W
    return out, kp


def tma_ok(x, P):
    """The TMA-fed TF32 kernel (csrc/tc_cm_gemm.cu) addresses rows of P floats: 16-byte aligned rows only."""
    return _PRECISION == "tf32" and P % 4 == 0 and x.data_ptr() % 16 == 0


def tiny(cin, cout, P):
    """Layers apb_pwconv_fwd_tf32 computes in exact fp32 on the CUDA cores (csrc/tc_cm_gemm.cu: <= 16 output or <= 8 input
    channels, <= 2048 weights, >= 256 points -- the imitator's mask head and the xyz embeddings)."""
    return (cout <= 16 or cin <= 8) and cin * cout <= 2048 and P >= 256


def _gemm_rows(x, w2d, bias, B, cin, cout, P, stats=None):
    """y (B,cout,P) = w2d (cout,cin) . x (B,cin,P) (+bias) through apb_pwconv_fwd_{tf32,bf16}.  The TMA-fed TF32 kernel
    takes any cout in one launch (and can accumulate the BatchNorm statistics of y into `stats` (cout,2) f64); the
    register-staged kernels are called in slices of <= 512 output channels."""
    if tma_ok(x, P) or (_PRECISION == "tf32" and tiny(cin, cout, P) and stats is None):
        wb, kp = pack_f32(w2d)
        y = torch.empty((B, cout, P), dtype=torch.float32, device=x.device)
        shift = None if bias is None else bias.contiguous().float()
        if stats is None:
            _cabi.call("apb_pwconv_fwd_tf32", [x, wb, None, shift, 0, y, B, cin, cout, P, kp], x.device)
        else:
            _cabi.call("apb_pwconv_ex_tf32", [x, wb, None, shift, 0, y, 0, stats, None, B, cin, cout, P, kp], x.device)
        return y
    if stats is not None:
        raise _cabi.ApbError("fused BatchNorm statistics need the TMA-fed kernel (P % 4 == 0)")
    outs = []
    entry = "apb_pwconv_fwd_tf32" if _PRECISION == "tf32" else "apb_pwconv_fwd_bf16"
    for c0 in range(0, cout, _MAX_N):
        c1 = min(cout, c0 + _MAX_N)
        wb, kp = (pack_f32 if _PRECISION == "tf32" else pack_bf16)(w2d[c0:c1])
        y = torch.empty((B, c1 - c0, P), dtype=torch.float32, device=x.device)
        if bias is not None:
            scale = torch.ones(c1 - c0, dtype=torch.float32, device=x.device)
            shift = bias[c0:c1].contiguous().float()
        else:
            scale = shift = None
        _cabi.call(entry, [x, wb, scale, shift, 0, y, B, cin, c1 - c0, P, kp], x.device)
        outs.append(y)
    return outs[0] if len(outs) == 1 else torch.cat(outs, dim=1)


def wgrad(g, x, B, cin, cout, P):
    """gW (cout, cin) = sum_{b,p} g[b,:,p] x[b,:,p]^T on the TMA-fed TF32 kernel; the accumulator rows are padded to a
    multiple of 4 floats so that the split-K partial sums land with 128-bit reductions."""
    ldw = (cin + 3) // 4 * 4
    gw = torch.zeros((cout, ldw), dtype=torch.float32, device=g.device)
    _cabi.call("apb_pwconv_wgrad_ld_tf32", [g, x, gw, ldw, B, cin, cout, P], g.device)
    return gw if ldw == cin else gw[:, :cin].contiguous()


class PointwiseConvFn(Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stats=None):
        """stats: optional pre-zeroed (Cout,2) f64 tensor that receives sum / sum of squares of the output (the batch
        statistics of a BatchNorm that follows), accumulated in the GEMM epilogue."""
        shp = x.shape
        B, cin = shp[0], shp[1]
        P = 1
        for d in shp[2:]:
            P *= d
        cout = weight.shape[0]
        xs = x.detach().contiguous().float()
        w2d = weight.detach().reshape(cout, cin).float()
        y = _gemm_rows(xs, w2d, None if bias is None else bias.detach(), B, cin, cout, P, stats)
        ctx.save_for_backward(xs, weight)
        ctx.has_bias = bias is not None
        ctx.dims = (B, cin, cout, P, tuple(shp))
        return y.view((B, cout) + tuple(shp[2:]))

    @staticmethod
    def backward(ctx, gy):
        xs, weight = ctx.saved_tensors
        B, cin, cout, P, shp = ctx.dims
        g = gy.detach().contiguous().float().view(B, cout, P)
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            wt = weight.detach().reshape(cout, cin).float().t().contiguous()
            gx = _gemm_rows(g, wt, None, B, cout, cin, P).view(shp)
        if ctx.needs_input_grad[1]:
            if P % 4 == 0:
                gw = wgrad(g, xs.view(B, cin, P), B, cin, cout, P)
            else:  # TMA needs 16-byte aligned rows; odd point counts do not occur on the PointNeXt path
                gw = torch.einsum("bop,bip->oi", g, xs.view(B, cin, P))
            gw = gw.view(weight.shape)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = g.sum(dim=(0, 2))
        return gx, gw, gb, None


class PointwiseConvMaxFn(Function):
    """pooled (B,Cout) = max over the points of (W x + b), x (B,Cin,P).  On the TMA-fed TF32 kernel the max and its
    first arg-max are taken in the GEMM epilogue (apb_pwconv_ex_tf32 with `pmax`): the (B,Cout,P) activation is never
    written.  Shapes TMA cannot address (P % 4 != 0) run the conv in slices of <= 512 output channels, each reduced by
    apb_rowmax_f32 and dropped.  The backward pass touches only the arg-max columns.  Callers apply ReLU to the pooled
    result (max(relu(y)) == relu(max(y)))."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        B, cin = x.shape[0], x.shape[1]
        P = x[0, 0].numel()
        cout = weight.shape[0]
        xs = x.detach().contiguous().float().view(B, cin, P)
        w2d = weight.detach().reshape(cout, cin).float()
        if tma_ok(xs, P) and not tiny(cin, cout, P):
            wb, kp = pack_f32(w2d)
            packed = torch.zeros((B, cout), dtype=torch.int64, device=x.device)
            shift = None if bias is None else bias.detach().contiguous().float()
            # (point_major = 2 is ignored when `pmax` is given; it marks the call for bench.py's byte model)
            _cabi.call("apb_pwconv_ex_tf32", [xs, wb, None, shift, 0, None, 2, None, packed, B, cin, cout, P, kp],
                       x.device)
            val = torch.empty((B, cout), dtype=torch.float32, device=x.device)
            arg32 = torch.empty((B, cout), dtype=torch.int32, device=x.device)
            _cabi.call("apb_pmax_decode", [packed, val, arg32, B * cout], x.device)
            ctx.save_for_backward(xs, weight, arg32.long())
            ctx.has_bias = bias is not None
            ctx.xshape = tuple(x.shape)
            return val
        vals, args = [], []
        for c0 in range(0, cout, _MAX_N):
            c1 = min(cout, c0 + _MAX_N)
            y = _gemm_rows(xs, w2d[c0:c1], None if bias is None else bias.detach()[c0:c1], B, cin, c1 - c0, P)
            v = torch.empty((B, c1 - c0), dtype=torch.float32, device=x.device)
            a = torch.empty((B, c1 - c0), dtype=torch.int32, device=x.device)
            _cabi.call("apb_rowmax_f32", [y, v, a, B * (c1 - c0), P], x.device)
            vals.append(v)
            args.append(a)
            del y
        val = vals[0] if len(vals) == 1 else torch.cat(vals, 1)
        arg = (args[0] if len(args) == 1 else torch.cat(args, 1)).long()
        ctx.save_for_backward(xs, weight, arg)
        ctx.has_bias = bias is not None
        ctx.xshape = tuple(x.shape)
        return val

    @staticmethod
    def backward(ctx, gy):
        xs, weight, arg = ctx.saved_tensors
        B, cin, P = xs.shape
        cout = weight.shape[0]
        g = gy.detach().float()
        idx = arg.unsqueeze(1).expand(B, cin, cout)
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            w2d = weight.detach().reshape(cout, cin).float()
            contrib = g.unsqueeze(1) * w2d.t().unsqueeze(0)                      # (B,cin,cout)
            gx = torch.zeros_like(xs).scatter_add_(2, idx, contrib).view(ctx.xshape)
        if ctx.needs_input_grad[1]:
            gw = torch.einsum("bo,bio->oi", g, xs.gather(2, idx)).view(weight.shape)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = g.sum(0)
        return gx, gw, gb


def conv_max(mod, x):
    """max over all positions of the 1x1 conv `mod` applied to x (B,Cin,...) -> (B,Cout); fused path on the tc
    backend, plain torch otherwise."""
    if _use_tc(mod, x):
        return PointwiseConvMaxFn.apply(x, mod.weight, mod.bias)
    y = mod(x)
    return y.flatten(2).max(dim=2)[0]


def _use_tc(mod, x):
    return (_BACKEND == "tc" and x.is_cuda and x.dtype == torch.float32 and mod.groups == 1
            and all(k == 1 for k in mod.kernel_size) and all(s == 1 for s in mod.stride)
            and all(p == 0 for p in mod.padding))


def conv_with_stats(mod, x):
    """(y, stats): y = mod(x) and, when the TMA-fed kernel can run this shape, stats (Cout,2) f64 = sum / sum of
    squares of y over (batch, positions) from the GEMM epilogue; stats is None otherwise (caller runs its own pass)."""
    if _use_tc(mod, x):
        P = x[0, 0].numel()
        if _PRECISION == "tf32" and P % 4 == 0 and not tiny(mod.in_channels, mod.out_channels, P):
            stats = torch.zeros((mod.out_channels, 2), dtype=torch.float64, device=x.device)
            return PointwiseConvFn.apply(x, mod.weight, mod.bias, stats), stats
    return mod(x), None


class TcConv1d(nn.Conv1d):
    def forward(self, x):
        if _use_tc(self, x):
            return PointwiseConvFn.apply(x, self.weight, self.bias)
        return super().forward(x)


class TcConv2d(nn.Conv2d):
    def forward(self, x):
        if _use_tc(self, x):
            return PointwiseConvFn.apply(x, self.weight, self.bias)
        return super().forward(x)
