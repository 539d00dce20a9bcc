"""Module/model parity on the GPU: our modules over the real sm_100a kernels (C-ABI) against the goldens the
reference's own Python produced (tests/golden/make_model_golden.py, make_step_golden.py).  Every case runs as
  strict : conv_backend "torch" (TEST-ONLY mode: cuDNN fp32 convs / BatchNorm over our ops), fp32 limits;
  tc     : the PRODUCT path -- the default backend: TF32 tcgen05 1x1 convs (fwd, dgrad, wgrad), our BatchNorm /
           BatchNorm->max / attention / deformation kernels.  Judged against the same fp32 goldens, relative to what the
           reference's own default precision (cuDNN with TF32 allowed) loses on the same fixture
           (_model_cases.check: value <= max(3 x cuDNN-TF32 deviation, TF32 floor))."""
import pytest
import torch

import _model_cases as mc

pytestmark = pytest.mark.gpu


def _show(tag, log):
    print(f"\n[{tag}] " + "; ".join(f"{n}={v:.2e}" + ("" if lim is None else f" (<= {lim:.2e})") for _, n, v, lim in log))


def _run_strict(fn):
    from adaptpoint_b200 import tcconv
    mc.strict_fp32()
    log = mc.begin("strict")
    try:
        with tcconv.conv_backend("torch"):
            fn("torch")
    finally:
        _show("strict", log)


def _run_tc_relative(fn):
    from adaptpoint_b200 import tcconv
    try:
        # (1) the reference's default precision: stock torch layers, cuDNN convolutions with TF32 allowed
        torch.backends.cudnn.allow_tf32 = True
        torch.backends.cuda.matmul.allow_tf32 = False
        log = mc.begin("record")
        with tcconv.conv_backend("torch"):
            fn("torch")
        _show("cuDNN-TF32", log)
        # (2) the product path
        mc.strict_fp32()
        log = mc.begin("relative")
        try:
            with tcconv.conv_backend("tc"):
                fn("tc")
        finally:
            _show("tc", log)
    finally:
        mc.strict_fp32()
        mc.begin("strict")


@pytest.mark.parametrize("profile", ["strict", "tc"])
@pytest.mark.parametrize("case", mc.ALL, ids=lambda f: f.__name__)
def test_module_parity_gpu(cuda_dev, case, profile):
    (_run_strict if profile == "strict" else _run_tc_relative)(lambda backend: case(cuda_dev))


@pytest.mark.parametrize("profile,kw", [("strict", {}), ("tc", {}), ("tc", {"fused_eval": True}), ("tc", {"overlap": False})],
                         ids=["strict", "tc", "tc-fused_eval", "tc-serial"])
def test_adaptpoint_iteration_vs_reference_gpu(cuda_dev, profile, kw):
    """One whole AdaptPoint iteration (train_gan body + train_one_epoch body) against the reference's own modules and
    feedback loss (tests/golden/model_step.npz): losses, generated cloud, logits and per-network gradients -- on the
    benchmarked configuration (tc backend, side-stream branches, optional fused bf16 no-grad classifier pass)."""
    if profile == "strict":
        _run_strict(lambda backend: mc.case_step(cuda_dev, backend, **kw))
    else:
        # the cuDNN-TF32 record run never uses the fused bf16 pass (it does not exist on the torch backend)
        _run_tc_relative(lambda backend: mc.case_step(cuda_dev, backend, **(kw if backend == "tc" else
                                                                             {k: v for k, v in kw.items() if k != "fused_eval"})))
