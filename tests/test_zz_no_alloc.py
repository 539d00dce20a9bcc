"""The no-allocation ball query (apb_ball_query_grid_static_f32, csrc/query_grid.cu,
no counterpart in the reference): workspace in a device array built into the library instead of caller memory.
Checked through the C-ABI against the existing entry point (apb_ball_query_grid_f32 with a caller workspace) -- bit-exact,
the tolerance of every ball-query test here -- on two calls in a row of different shapes (the second call finds the first
call's contents in the shared array), plus the size limit and the absence of any device allocation.
Named test_zz_* so that pytest collects it last."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _query(name, dev, q, xyz, B, N, M, radius, K, ws=None):
    from adaptpoint_b200 import _cabi
    idx = torch.full((B, M, K), -7, dtype=torch.int32, device=dev)
    if ws is None:
        _cabi.call(name, [q, xyz, idx, B, N, M, float(radius), K], dev)
    else:
        _cabi.call(name, [q, xyz, idx, ws, ws.numel(), B, N, M, float(radius), K], dev)
    return idx


def test_ball_query_static_scratch_matches_workspace_entry_point(cuda_dev):
    from adaptpoint_b200 import _cabi
    from adaptpoint_b200.synth import make_clouds
    lib = _cabi.load()
    limit = lib.apb_ball_query_grid_static_ws_bytes()
    assert limit == 16 << 20
    # two calls in a row: a large cloud first, then a smaller batch of a different kind whose cell tables land on the
    # first call's leftovers
    cases = [(8, 24000, 6000, 0.1, 32, "ball", 1), (3, 5000, 1250, 0.05, 16, "shell", 2)]
    prepared = []
    for B, N, M, radius, K, kind, seed in cases:
        xyz = make_clouds(B, N, kind, seed).to(cuda_dev)
        q = xyz[:, torch.randperm(N, generator=torch.Generator().manual_seed(seed))[:M]].contiguous()
        need = lib.apb_ball_query_grid_ws_bytes(B, N)
        assert 0 < need <= limit
        ws = torch.empty(need, dtype=torch.uint8, device=cuda_dev)
        want = _query("apb_ball_query_grid_f32", cuda_dev, q, xyz, B, N, M, radius, K, ws)
        prepared.append((B, N, M, radius, K, xyz, q, want))
    torch.cuda.synchronize()
    for rep in range(2):          # the pair twice: every call after the first also checks "no allocation"
        for B, N, M, radius, K, xyz, q, want in prepared:
            idx = torch.full((B, M, K), -7, dtype=torch.int32, device=cuda_dev)
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(cuda_dev)[0]
            _cabi.call("apb_ball_query_grid_static_f32", [q, xyz, idx, B, N, M, float(radius), K], cuda_dev)
            torch.cuda.synchronize()
            free_after = torch.cuda.mem_get_info(cuda_dev)[0]
            assert torch.equal(idx, want), f"rep {rep}, B={B} N={N}"
            if rep == 1:          # (the very first call may load the kernels' module lazily)
                assert free_after == free_before, "the entry point must not allocate device memory"


def test_ball_query_static_scratch_rejects_oversize(cuda_dev):
    from adaptpoint_b200 import _cabi
    lib = _cabi.load()
    B, N, M, K = 64, 100000, 8, 16
    assert lib.apb_ball_query_grid_ws_bytes(B, N) > lib.apb_ball_query_grid_static_ws_bytes()
    xyz = torch.zeros((1, 8, 3), device=cuda_dev)        # never dereferenced: the size check comes first
    idx = torch.zeros((1, M, K), dtype=torch.int32, device=cuda_dev)
    with pytest.raises(_cabi.ApbError, match="built-in scratch"):
        _cabi.call("apb_ball_query_grid_static_f32", [xyz, xyz, idx, B, N, M, 0.1, K], cuda_dev)
    status = lib.apb_ball_query_grid_static_f32(ctypes.c_void_p(xyz.data_ptr()), ctypes.c_void_p(xyz.data_ptr()),
                                                ctypes.c_void_p(idx.data_ptr()), B, N, M, 0.1, K, None)
    assert status == 10003        # APB_E_SCRATCH_LIMIT
