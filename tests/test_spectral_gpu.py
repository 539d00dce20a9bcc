"""Fused spectral normalisation (spectral.py, csrc/spectral.cu) against torch.nn.utils.parametrizations.spectral_norm:
same normalised weight, same in-place power-iteration update of u / v, same gradient, same state_dict layout."""
import copy

import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30)).item()


@pytest.mark.parametrize("make", [lambda: nn.Linear(1024, 512), lambda: nn.Linear(15, 1), lambda: nn.Conv2d(3, 64, 1),
                                  lambda: nn.Conv2d(128, 1024, 1), lambda: nn.Linear(256, 15)],
                         ids=["fc1", "prob_head", "conv1", "conv3", "fc3"])
def test_fused_spectral_norm_matches_torch(cuda_dev, make):
    from adaptpoint_b200 import spectral, tcconv
    torch.manual_seed(3)
    base = make()
    ref = nn.utils.parametrizations.spectral_norm(copy.deepcopy(base)).to(cuda_dev).train()
    ours = spectral.spectral_norm(copy.deepcopy(base)).to(cuda_dev).train()
    assert list(ours.state_dict().keys()) == list(ref.state_dict().keys())
    ours.load_state_dict(ref.state_dict())
    # (base.weight, not ref.weight: reading a parametrized weight in training mode runs a power iteration)
    g = torch.randn(base.weight.shape, generator=torch.Generator().manual_seed(1)).to(cuda_dev)
    with tcconv.conv_backend("tc"):
        for step in range(3):                        # three forwards: u, v keep evolving in place exactly like torch's
            w_ref, w_ours = ref.weight, ours.weight
            assert _rel(w_ours, w_ref) < 2e-5, step
        (w_ref * g).sum().backward()
        (w_ours * g).sum().backward()
    for k in ("_u", "_v"):
        a, b = getattr(ours.parametrizations.weight[0], k), getattr(ref.parametrizations.weight[0], k)
        assert _rel(a, b) < 2e-5, k
    ga = ours.parametrizations.weight.original.grad
    gb = ref.parametrizations.weight.original.grad
    assert _rel(ga, gb) < 5e-5
    # eval mode takes torch's own path (no power iteration): identical by construction
    ours.eval(); ref.eval()
    assert _rel(ours.weight, ref.weight) < 2e-5
