"""Spectral normalisation on libapb200 (csrc/spectral.cu): 3 launches forward, 2 backward per layer.

Drop-in for torch.nn.utils.parametrizations.spectral_norm as PointDiscriminator1 uses it
(openpoints/models_adaptpoint/point_discriminator.py:23-34): same parametrization class hierarchy, same state_dict entries
(`parametrizations.weight.original`, `parametrizations.weight.0._u`, `._v`), same semantics -- one power iteration per
training-mode forward that updates u, v in place, sigma = u^T W v with u, v treated as constants by autograd.  On CUDA
fp32 weights in training mode the arithmetic runs in `SpectralNormFn`; everything else (CPU tensors of the op-patched
tests, eval mode, exotic settings) takes torch's own implementation.
"""
import torch
from torch.autograd import Function
from torch.nn.utils import parametrize
from torch.nn.utils.parametrizations import _SpectralNorm

from . import _cabi, tcconv


class SpectralNormFn(Function):
    @staticmethod
    def forward(ctx, weight, u, v, eps):
        h = weight.shape[0]
        wm = weight.detach().contiguous().view(h, -1)
        w = wm.shape[1]
        dev = wm.device
        scratch = torch.zeros(4 + w, dtype=torch.float32, device=dev)
        t_ws = torch.empty(h, dtype=torch.float32, device=dev)
        us, vs = torch.empty(h, dtype=torch.float32, device=dev), torch.empty(w, dtype=torch.float32, device=dev)
        sigma = torch.empty(1, dtype=torch.float32, device=dev)
        out = torch.empty_like(wm)
        _cabi.call("apb_spectral_norm_fwd_f32", [wm, u, v, t_ws, scratch, us, vs, sigma, out, float(eps), h, w], dev)
        ctx.save_for_backward(out, us, vs, sigma)
        return out.view(weight.shape)

    @staticmethod
    def backward(ctx, g):
        out, us, vs, sigma = ctx.saved_tensors
        h, w = out.shape
        gm = g.detach().contiguous().float().view(h, w)
        scratch = torch.zeros(4, dtype=torch.float32, device=gm.device)
        dw = torch.empty_like(out)
        _cabi.call("apb_spectral_norm_bwd_f32", [gm, out, us, vs, sigma, scratch, dw, h, w], gm.device)
        return dw.view(g.shape), None, None, None


class FusedSpectralNorm(_SpectralNorm):
    def forward(self, weight):
        if (weight.is_cuda and weight.dtype == torch.float32 and weight.ndim > 1 and self.training and self.dim == 0
                and self.n_power_iterations == 1 and tcconv.get_conv_backend() == "tc"):
            return SpectralNormFn.apply(weight, self._u, self._v, self.eps)
        return super().forward(weight)


def spectral_norm(module, name="weight", n_power_iterations=1, eps=1e-12):
    """torch.nn.utils.parametrizations.spectral_norm(module, name) with the fused CUDA path above."""
    weight = getattr(module, name)
    parametrize.register_parametrization(module, name, FusedSpectralNorm(weight, n_power_iterations, 0, eps))
    return module
