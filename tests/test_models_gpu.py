"""Module/model parity on the GPU: our modules over the real sm_100a kernels (C-ABI) against the goldens the
reference's own Python produced (tests/golden/make_model_golden.py)."""
import pytest

import _model_cases as mc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", mc.ALL, ids=lambda f: f.__name__)
def test_module_parity_gpu(cuda_dev, case):
    mc.strict_fp32()
    case(cuda_dev)
