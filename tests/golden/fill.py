"""Deterministic parameter fill shared by make_model_golden.py (reference modules) and the parity tests (our
modules): every state_dict entry is generated from a CPU generator seeded by crc32(entry name), so equal names
and shapes get equal values -- which also checks state_dict compatibility with the reference."""
import zlib

import torch


def fill_state(model, seed=0):
    sd = model.state_dict()
    new = {}
    for name, t in sd.items():
        g = torch.Generator(device="cpu").manual_seed((zlib.crc32(name.encode()) + seed) % (2 ** 31))
        shape = tuple(t.shape)
        u = lambda lo, hi: torch.rand(shape, generator=g) * (hi - lo) + lo
        if name.endswith("num_batches_tracked"):
            v = torch.zeros(shape, dtype=t.dtype)
        elif name.endswith("running_var") or "affine_alpha" in name:
            v = u(0.5, 1.5)
        elif name.endswith("running_mean") or "affine_beta" in name:
            v = u(-0.2, 0.2)
        elif name.endswith("_u") or name.endswith("_v"):
            v = torch.randn(shape, generator=g)
            v = v / v.norm()
        elif t.dim() >= 2:
            fan_in = 1
            for s in shape[1:]:
                fan_in *= s
            a = (3.0 / fan_in) ** 0.5
            v = u(-a, a)
        elif name.endswith("weight"):
            v = u(0.5, 1.5)
        else:
            v = u(-0.1, 0.1)
        new[name] = v.to(t.dtype)
    model.load_state_dict(new, strict=True)
    return sorted((n, tuple(v.shape)) for n, v in new.items())
