"""Host logic of the module layer (layers.py, pointnext.py, models.py, imitator.py, train.py) against the goldens
produced by the reference's own Python (tests/golden/make_model_golden.py).  No GPU here: adaptpoint_b200.ops is
pointed at the CPU stand-ins of oracle/torch_ops.py for the duration of each test, so what is checked is OUR module
wiring, parameter naming and math -- the CUDA kernels are checked by the -m gpu suites."""
import pytest
import torch

import _model_cases as mc
from oracle import torch_ops


@pytest.mark.parametrize("case", mc.ALL, ids=lambda f: f.__name__)
def test_module_parity_cpu(case):
    with torch_ops.patched():
        mc.begin("strict")
        case("cpu")


def test_adaptpoint_iteration_vs_reference_cpu():
    """One whole AdaptPoint iteration (imitator + discriminator + classifier updates) of OUR trainer on the CPU stand-in
    ops against the reference's own modules / feedback loss (tests/golden/model_step.npz)."""
    with torch_ops.patched():
        mc.begin("strict")
        mc.case_step("cpu", "torch")


def test_trainer_step_cpu():
    from adaptpoint_b200.synth import make_clouds
    with torch_ops.patched():
        from adaptpoint_b200.train import AdaptPointTrainer
        tr = AdaptPointTrainer(device="cpu")
        B, N = 4, 512
        pos = make_clouds(B, N, "shell", 0)
        pts = torch.cat([pos, pos[:, :, 1:2] - pos[:, :, 1:2].amin(1, keepdim=True)], -1)
        y = torch.randint(0, 15, (B,), generator=torch.Generator().manual_seed(0))
        before = [p.detach().clone() for p in tr.generator.parameters()][:3]
        losses = tr.step(pts, y)
        assert all(torch.isfinite(l) for l in losses)
        after = list(tr.generator.parameters())[:3]
        assert any((a - b).abs().max() > 0 for a, b in zip(after, before)), "generator was not updated"


def test_product_ops_refuse_cpu_tensors():
    from adaptpoint_b200 import ops, _cabi
    with pytest.raises(_cabi.ApbError):
        ops.furthest_point_sample(torch.zeros(1, 8, 3), 4)
