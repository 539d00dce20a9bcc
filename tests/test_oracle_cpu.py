"""CPU suite (-m "not gpu"): pins the oracle (against an independent numpy statement of the
reference's tie rule, hand-computed known answers and the committed golden vectors), checks the
host-side logic, and that libapb200.so loads and exports every symbol include/apb200.h declares."""
import ctypes
import glob
import os
import re

import numpy as np
import pytest

from adaptpoint_b200.synth import make_clouds
from oracle import cpu as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _sqd(a, b):
    """reference contraction order (see oracle/pointnet2_oracle.c sqdist_ref) in numpy float32/float64:
    t = dy*dy (rounded) ; fma(dx,dx,t) ; fma(dz,dz,t) -- emulated exactly through float64 products
    (a float32 product is exact in float64; one rounding of the sum = fused)."""
    d = (a - b).astype(np.float32)
    t = (d[..., 1] * d[..., 1]).astype(np.float32)
    t = (d[..., 0].astype(np.float64) * d[..., 0].astype(np.float64) + t.astype(np.float64)).astype(np.float32)
    t = (d[..., 2].astype(np.float64) * d[..., 2].astype(np.float64) + t.astype(np.float64)).astype(np.float32)
    return t


def _bitrev(v, bits):
    r = 0
    for _ in range(bits):
        r = (r << 1) | (v & 1)
        v >>= 1
    return r


def fps_numpy_total_order(xyz, m):
    """Independent statement: argmax of min-dist under the total order
    (value desc, bitrev(k mod bs) asc, k asc) -- SURVEY.md 8a -- instead of emulating threads."""
    N = xyz.shape[0]
    bs = orc.opt_n_threads(N)
    L = bs.bit_length() - 1
    k = np.arange(N)
    key2 = np.array([_bitrev(int(v) % bs, L) for v in k])
    md = np.full(N, 1e10, np.float32)
    out = [0]
    for _ in range(1, m):
        md = np.minimum(_sqd(xyz, xyz[out[-1]]), md)
        best = md.max()
        cand = k[md == best]
        cand = sorted(cand, key=lambda c: (key2[c], c))
        out.append(int(cand[0]))
    return np.array(out, np.int32)


@pytest.mark.parametrize("N,m,kind", [(64, 32, "quant"), (100, 100, "same"), (300, 120, "quant"), (1024, 200, "quant"),
                                      (1200, 150, "dup"), (2048, 64, "quant"), (31, 31, "quant"), (1500, 40, "ball")])
def test_oracle_fps_matches_total_order_statement(N, m, kind):
    xyz = make_clouds(2, N, kind, seed=N).numpy()
    got = orc.fps(xyz, m)
    for b in range(2):
        np.testing.assert_array_equal(got[b], fps_numpy_total_order(xyz[b], m))


def test_fps_known_answer_tie_rule():
    """SURVEY.md 8a worked example: N = bs = 1024, points 1 and 2 equally far from point 0 and all
    others closer -> the reference picks 2 (bit-reversed thread id 256 < 512), not the first index."""
    xyz = np.zeros((1, 1024, 3), np.float32)
    xyz[0, 1] = (1, 0, 0)
    xyz[0, 2] = (0, 1, 0)
    assert orc.fps(xyz, 2)[0].tolist() == [0, 2]
    # all points identical: every round is an all-way tie at value 0 -> slot 0 (k=0) wins
    same = np.full((1, 50, 3), 0.5, np.float32)
    assert orc.fps(same, 4)[0].tolist() == [0, 0, 0, 0]
    # 1-D line: farthest from 0 is the far end, then the middle
    line = np.zeros((1, 5, 3), np.float32)
    line[0, :, 0] = [0, 1, 2, 3, 4]
    assert orc.fps(line, 3)[0].tolist() == [0, 4, 2]


def test_ball_query_known_answers():
    xyz = np.zeros((1, 6, 3), np.float32)
    xyz[0, :, 0] = [0, 0.5, 1.0, 1.5, 2.0, 0.25]
    q = np.zeros((1, 2, 3), np.float32)
    q[0, 1, 0] = 10.0
    idx = orc.ball_query(1.0, 4, xyz, q)
    # strict '<': the point at distance exactly 1.0 is outside; first hit pads the row; no hit -> zeros
    assert idx[0, 0].tolist() == [0, 1, 5, 0]
    assert idx[0, 1].tolist() == [0, 0, 0, 0]


def test_three_nn_known_answers():
    known = np.zeros((1, 4, 3), np.float32)
    known[0, :, 0] = [3, 1, 1, 2]
    unk = np.zeros((1, 1, 3), np.float32)
    d2, idx = orc.three_nn(unk, known)
    assert idx[0, 0].tolist() == [1, 2, 3]  # ties keep index order
    assert d2[0, 0].tolist() == [1.0, 1.0, 4.0]
    d2, idx = orc.three_nn(unk, known[:, :2])  # m < 3: unfilled slot stays (inf, 0)
    assert idx[0, 0].tolist() == [1, 0, 0] and np.isinf(d2[0, 0, 2])


def test_oracle_group_gather_interp_against_numpy():
    rng = np.random.default_rng(0)
    feat = rng.standard_normal((2, 5, 40)).astype(np.float32)
    idx = rng.integers(0, 40, (2, 7, 3)).astype(np.int32)
    want = np.stack([feat[b][:, idx[b]] for b in range(2)])
    np.testing.assert_array_equal(orc.group(feat, idx), want)
    go = rng.standard_normal((2, 5, 7, 3)).astype(np.float32)
    wg = np.zeros((2, 5, 40), np.float64)
    for b in range(2):
        for c in range(5):
            np.add.at(wg[b, c], idx[b].ravel(), go[b, c].ravel().astype(np.float64))
    np.testing.assert_allclose(orc.group_grad(go, idx, 40), wg, rtol=1e-6, atol=1e-6)
    w = rng.random((2, 7, 3)).astype(np.float32)
    out = orc.three_interp(feat, idx, w)
    ref = np.einsum("bcnj,bnj->bcn", want.astype(np.float64), w.astype(np.float64))
    np.testing.assert_allclose(out, ref, rtol=1e-5, atol=1e-6)


def test_golden_vectors_pin_the_oracle():
    """tests/golden/*.npz were produced on the B200 box by tests/golden/make_golden.py from the
    reference's own CUDA kernels (oracle/_ref); the CPU restatement must reproduce them bit for bit."""
    files = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "case*.npz")))
    if not files:
        pytest.skip("golden vectors not generated yet")
    for f in files:
        z = np.load(f)
        xyz = z["xyz"]
        np.testing.assert_array_equal(orc.fps(xyz, int(z["fps_m"])), z["fps_idx"])
        np.testing.assert_array_equal(orc.ball_query(float(z["bq_radius"]), int(z["bq_k"]), xyz, z["bq_query"]), z["bq_idx"])
        d2, ix = orc.three_nn(xyz, z["bq_query"])
        np.testing.assert_array_equal(ix, z["nn_idx"])
        np.testing.assert_array_equal(d2, z["nn_d2"])
        np.testing.assert_array_equal(orc.group(z["feat"], z["bq_idx"]), z["group_out"])
        np.testing.assert_array_equal(orc.three_interp(z["feat_known"], z["nn_idx"], z["interp_w"]), z["interp_out"])


def test_cabi_library_loads_and_exports_header_symbols():
    hdr = open(os.path.join(ROOT, "include", "apb200.h")).read()
    declared = sorted(set(re.findall(r"\b(apb_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 13
    from adaptpoint_b200 import _cabi
    assert os.path.exists(_cabi.lib_path()), "libapb200.so not built (run __graft_entry__.build())"
    lib = ctypes.CDLL(_cabi.lib_path())
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/apb200.h but not exported"
    assert set(_cabi.EXPORTED_SYMBOLS) <= set(declared)
    assert "sm_100a" in _cabi.version()


def test_host_block_size_rule_matches_reference_rule():
    from adaptpoint_b200 import _cabi
    for n in list(range(1, 70)) + [127, 128, 129, 255, 256, 511, 512, 513, 1000, 1023, 1024, 1025, 2048, 4096, 8192,
                                   16384, 24000, 65536, 100000, 1 << 20]:
        assert _cabi.fps_block_size(n) == orc.opt_n_threads(n)


def test_product_never_imports_the_oracle():
    for path in glob.glob(os.path.join(ROOT, "adaptpoint_b200", "**", "*"), recursive=True):
        if path.endswith((".py", ".cu", ".cuh", ".h")):
            assert "oracle" not in open(path).read().replace("oracle/", "").lower() or path.endswith("_none"), path


def test_cpu_tensors_are_rejected_not_silently_computed():
    import torch
    from adaptpoint_b200 import _cabi, ops
    with pytest.raises(_cabi.ApbError):
        ops.ball_query(0.1, 4, torch.zeros(1, 8, 3), torch.zeros(1, 2, 3))
