"""Module/model parity on the GPU: our modules over the real sm_100a kernels (C-ABI) against the goldens the
reference's own Python produced (tests/golden/make_model_golden.py, make_step_golden.py).  Every case runs twice:
  strict : conv_backend "torch" (TEST-ONLY mode: cuDNN fp32 convs / BatchNorm over our ops), fp32 tolerances;
  tc     : the PRODUCT path -- the default backend: TF32 tcgen05 1x1 convs (fwd, dgrad, wgrad), our BatchNorm /
           BatchNorm->max / attention / deformation kernels -- at TF32 tolerances (_model_cases.PROFILES)."""
import pytest

import _model_cases as mc

pytestmark = pytest.mark.gpu


@pytest.fixture
def profile(request):
    from adaptpoint_b200 import tcconv
    prev = tcconv.get_conv_backend()
    tcconv.set_conv_backend("tc" if request.param == "tc" else "torch")
    mc.use_profile(request.param)
    try:
        yield request.param
    finally:
        tcconv.set_conv_backend(prev)
        mc.use_profile("strict")


@pytest.mark.parametrize("profile", ["strict", "tc"], indirect=True)
@pytest.mark.parametrize("case", mc.ALL, ids=lambda f: f.__name__)
def test_module_parity_gpu(cuda_dev, case, profile):
    mc.strict_fp32()
    case(cuda_dev)


@pytest.mark.parametrize("mode,kw", [("strict", {}), ("tc", {}), ("tc", {"fused_eval": True}), ("tc", {"overlap": False})],
                         ids=["strict", "tc", "tc-fused_eval", "tc-serial"])
def test_adaptpoint_iteration_vs_reference_gpu(cuda_dev, mode, kw):
    """One whole AdaptPoint iteration (train_gan body + train_one_epoch body) against the reference's own modules and
    feedback loss (tests/golden/model_step.npz): losses, generated cloud, logits and per-network gradients -- on the
    benchmarked configuration (tc backend, side-stream branches, optional fused bf16 no-grad classifier pass)."""
    mc.strict_fp32()
    mc.case_step(cuda_dev, mode, **kw)
