"""GPU parity tests: the sm_100a kernels (called through the C-ABI via adaptpoint_b200.ops) against
 (1) the CPU oracle (oracle/pointnet2_oracle.c) and
 (2) the reference's own CUDA kernels compiled unmodified (oracle/_ref/libpointnet2_ref.so).
Bar: bit-exact for indices and for forward copies / interpolation; 1e-5 relative for the
atomically-accumulated gradients (the reference itself sums them in a nondeterministic order)."""
import numpy as np
import pytest
import torch

from adaptpoint_b200 import ops
from adaptpoint_b200.synth import make_clouds
from oracle import cpu as orc
from oracle import refcuda

pytestmark = pytest.mark.gpu

REL_TOL = 1e-5  # north_star: fp32 features/gradients within 1e-5 relative


def _dev(t, dev):
    return t.to(dev).contiguous()


def _assert_close_rel(a, b, tol=REL_TOL):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    scale = max(np.abs(b).max(), 1e-30)
    err = np.abs(a - b).max() / scale
    assert err <= tol, f"max rel err {err:.3e} > {tol}"


FPS_CASES = [
    # B, N, m, kind
    (2, 1024, 512, "ball"), (2, 1024, 512, "quant"), (2, 2048, 1200, "shell"), (2, 2048, 4, "ball"),
    (3, 31, 7, "quant"), (2, 1200, 300, "dup"), (1, 512, 128, "quant"), (2, 256, 64, "quant"),
    (2, 128, 64, "dup"), (2, 100, 100, "same"), (2, 64, 32, "quant"), (2, 5, 5, "ball"), (1, 1, 1, "ball"),
    (2, 1024, 1, "ball"), (1, 700, 700, "quant"), (2, 1023, 256, "quant"), (2, 1025, 256, "quant"),
    (2, 4096, 1024, "quant"), (2, 4097, 500, "quant"), (2, 8192, 512, "quant"), (1, 16384, 1024, "quant"),
    (1, 24000, 600, "shell"), (1, 33000, 300, "quant"), (1, 65536, 200, "quant"), (1, 100000, 300, "shell"),
]


@pytest.mark.parametrize("B,N,m,kind", FPS_CASES)
def test_fps_vs_oracle(cuda_dev, B, N, m, kind):
    xyz = make_clouds(B, N, kind, seed=N + m)
    want = orc.fps(xyz.numpy(), m)
    got = ops.furthest_point_sample(_dev(xyz, cuda_dev), m).cpu().numpy()
    assert got.dtype == np.int32
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("B,N,m,kind", [
    (32, 1024, 512, "shell"), (32, 2048, 1200, "quant"), (32, 128, 64, "quant"), (8, 16384, 4096, "quant"),
    (4, 24000, 6000, "shell"), (2, 65536, 4096, "quant"), (2, 100000, 5000, "dup"), (40, 8192, 2048, "ball"),
])
def test_fps_vs_compiled_reference(cuda_dev, B, N, m, kind):
    if not refcuda.available():
        pytest.skip("oracle/_ref not built")
    xyz = _dev(make_clouds(B, N, kind, seed=7), cuda_dev)
    want = refcuda.fps(xyz, m)
    got = ops.furthest_point_sample(xyz, m)
    assert torch.equal(got, want)


def test_fps_temp_in_out(cuda_dev):
    """C-ABI extra: a caller-provided temp is the initial min-distance and receives the final one,
    exactly like the reference wrapper's scratch buffer."""
    from adaptpoint_b200 import _cabi
    xyz = _dev(make_clouds(2, 1500, "ball", 3), cuda_dev)
    temp = torch.full((2, 1500), 1e10, device=cuda_dev)
    idx = torch.empty((2, 100), dtype=torch.int32, device=cuda_dev)
    _cabi.call("apb_fps_f32", [xyz, temp, idx, 2, 1500, 100], cuda_dev)
    tnp = np.full((2, 1500), 1e10, np.float32)
    want = orc.fps(xyz.cpu().numpy(), 100, temp=tnp)
    np.testing.assert_array_equal(idx.cpu().numpy(), want)
    np.testing.assert_array_equal(temp.cpu().numpy(), tnp)


BQ_CASES = [
    # B, N, M, radius, K, kind
    (2, 1024, 512, 0.15, 32, "shell"), (2, 2048, 1024, 0.1, 24, "ball"), (2, 512, 256, 0.225, 32, "quant"),
    (2, 128, 64, 0.50625, 32, "quant"), (1, 5000, 700, 0.1, 32, "shell"), (2, 300, 300, 1e-4, 16, "ball"),
    (2, 20, 10, 0.5, 32, "ball"), (2, 1000, 130, 0.25, 64, "quant"), (1, 4500, 129, 0.125, 7, "quant"),
    (2, 256, 256, 10.0, 32, "dup"), (1, 100, 50, 0.3, 1, "same"),
]


def _queries(xyz, M):
    g = torch.Generator().manual_seed(M)
    sel = torch.stack([torch.randperm(xyz.shape[1], generator=g)[:M] for _ in range(xyz.shape[0])])
    return torch.gather(xyz, 1, sel.unsqueeze(-1).expand(-1, -1, 3)).contiguous()


@pytest.mark.parametrize("B,N,M,radius,K,kind", BQ_CASES)
def test_ball_query_vs_oracle_and_reference(cuda_dev, B, N, M, radius, K, kind):
    xyz = make_clouds(B, N, kind, seed=N)
    q = _queries(xyz, M)
    want = orc.ball_query(radius, K, xyz.numpy(), q.numpy())
    got = ops.ball_query(radius, K, _dev(xyz, cuda_dev), _dev(q, cuda_dev))
    np.testing.assert_array_equal(got.cpu().numpy(), want)
    if refcuda.available():
        ref = refcuda.ball_query(radius, K, _dev(xyz, cuda_dev), _dev(q, cuda_dev))
        assert torch.equal(got, ref)


@pytest.mark.parametrize("B,N,M,radius,K", [(32, 16384, 4096, 0.1, 32), (8, 24000, 6000, 0.1, 32), (4, 6000, 6000, 0.2, 32)])
def test_ball_query_full_size_vs_reference(cuda_dev, B, N, M, radius, K):
    if not refcuda.available():
        pytest.skip("oracle/_ref not built")
    xyz = _dev(make_clouds(B, N, "shell", seed=1), cuda_dev)
    q = xyz[:, :M].contiguous()
    assert torch.equal(ops.ball_query(radius, K, xyz, q), refcuda.ball_query(radius, K, xyz, q))


GROUP_CASES = [(2, 35, 1024, 512, 32), (2, 3, 2048, 1024, 24), (1, 67, 512, 256, 32), (2, 5, 100, 33, 7),
               (1, 259, 128, 64, 32), (1, 2, 60000, 100, 8), (2, 1, 17, 5, 3), (1, 6, 999, 250, 20),
               # rows longer than one SM's shared memory: sliced kernels with 2 / 3 / 5 slices (aligned and unaligned)
               (2, 3, 100000, 25000, 32), (1, 3, 120001, 40000, 7), (1, 2, 250000, 32768, 16), (1, 2, 60000, 20000, 8)]


@pytest.mark.parametrize("B,C,N,M,K", GROUP_CASES)
def test_group_fwd_bwd(cuda_dev, B, C, N, M, K):
    g = torch.Generator().manual_seed(C * N)
    feat = torch.randn(B, C, N, generator=g)
    idx = torch.randint(0, N, (B, M, K), generator=g, dtype=torch.int32)
    want = orc.group(feat.numpy(), idx.numpy())
    fd = _dev(feat, cuda_dev).requires_grad_(True)
    idd = _dev(idx, cuda_dev)
    got = ops.grouping_operation(fd, idd)
    np.testing.assert_array_equal(got.detach().cpu().numpy(), want)
    go = torch.randn(B, C, M, K, generator=g)
    got.backward(_dev(go, cuda_dev))
    want_g = orc.group_grad(go.numpy(), idx.numpy(), N)
    _assert_close_rel(fd.grad.cpu().numpy(), want_g)
    if refcuda.available():
        assert torch.equal(got.detach(), refcuda.group(fd.detach(), idd))
        _assert_close_rel(refcuda.group_grad(_dev(go, cuda_dev), idd, N).cpu().numpy(), want_g)


@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("B,C,N,M,K", [(2, 35, 1024, 512, 32), (2, 5, 101, 33, 7), (1, 3, 16384, 4096, 32), (1, 9, 30000, 700, 16),
                                       (1, 2, 70001, 30000, 6)])
def test_group_grad_both_paths_accumulate(cuda_dev, B, C, N, M, K, mode):
    """Both backward schemes (scattered L2 reductions = the reference's, group_points_gpu.cu:14-31; accumulator rows in
    shared memory) ADD into the caller's buffer, as the reference's atomicAdd does; tolerance 1e-5 relative."""
    from adaptpoint_b200 import _cabi
    g = torch.Generator().manual_seed(N + mode)
    idx = torch.randint(0, N, (B, M, K), generator=g, dtype=torch.int32)
    idx[:, :, K // 2:] = idx[:, :, :1]          # ball-query style padding: heavy duplicates inside a row
    go = torch.randn(B, C, M, K, generator=g)
    base = torch.randn(B, C, N, generator=g)
    want = base.numpy() + orc.group_grad(go.numpy(), idx.numpy(), N)
    acc = _dev(base, cuda_dev).clone()
    _cabi.tune_group_grad(mode)
    try:
        _cabi.call("apb_group_grad_f32", [_dev(go, cuda_dev), _dev(idx, cuda_dev), acc, B, C, N, M, K], acc.device)
    finally:
        _cabi.tune_group_grad(0)
    _assert_close_rel(acc.cpu().numpy(), want)


def test_group_grad_full_size_conserves_mass(cuda_dev):
    """cfg3 size (B=32, N=16384, M=4096, K=32, C=32): per (cloud, channel) the scattered gradient sums to the sum of
    grad_out (size-independent property; the oracle would take minutes here)."""
    B, C, N, M, K = 32, 32, 16384, 4096, 32
    g = torch.Generator(device=cuda_dev).manual_seed(3)
    idx = torch.randint(0, N, (B, M, K), generator=g, dtype=torch.int32, device=cuda_dev)
    go = torch.rand(B, C, M, K, generator=g, device=cuda_dev)
    feat = torch.zeros(B, C, N, device=cuda_dev, requires_grad=True)
    ops.grouping_operation(feat, idx).backward(go)
    got, want = feat.grad.double().sum(-1), go.double().sum((-1, -2))
    assert torch.allclose(got, want, rtol=1e-5)
    hist = torch.zeros(B, N, device=cuda_dev).scatter_add_(1, idx.view(B, -1).long(), go[:, 0].reshape(B, -1))
    assert torch.allclose(feat.grad[:, 0], hist, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("B,C,N,m", [(2, 128, 2048, 512), (2, 3, 10000, 4096), (1, 7, 33, 33), (2, 64, 1024, 1)])
def test_gather_fwd_bwd(cuda_dev, B, C, N, m):
    g = torch.Generator().manual_seed(C + N)
    feat = torch.randn(B, C, N, generator=g)
    idx = torch.randint(0, N, (B, m), generator=g, dtype=torch.int32)
    fd = _dev(feat, cuda_dev).requires_grad_(True)
    got = ops.gather_operation(fd, _dev(idx, cuda_dev))
    np.testing.assert_array_equal(got.detach().cpu().numpy(), orc.gather(feat.numpy(), idx.numpy()))
    # the reference's only self-check (subsample.py:159-185): gather_operation == torch.gather
    tg = torch.gather(fd.detach(), 2, _dev(idx, cuda_dev).long().unsqueeze(1).expand(-1, C, -1))
    assert torch.equal(got.detach(), tg)
    go = torch.randn(B, C, m, generator=g)
    got.backward(_dev(go, cuda_dev))
    _assert_close_rel(fd.grad.cpu().numpy(), orc.gather_grad(go.numpy(), idx.numpy(), N))


NN_CASES = [(2, 375, 93, "shell"), (2, 1500, 375, "ball"), (1, 6000, 1500, "quant"), (2, 256, 128, "quant"),
            (2, 50, 2, "ball"), (2, 50, 1, "ball"), (1, 3000, 2500, "dup"), (2, 64, 64, "same")]


@pytest.mark.parametrize("B,n,m,kind", NN_CASES)
def test_three_nn_and_interpolate(cuda_dev, B, n, m, kind):
    pts = make_clouds(B, max(n, m), kind, seed=n + m)
    unknown = pts[:, :n].contiguous()
    known = make_clouds(B, m, kind, seed=n) if kind not in ("same",) else pts[:, :m].contiguous()
    want_d2, want_idx = orc.three_nn(unknown.numpy(), known.numpy())
    ud, kd = _dev(unknown, cuda_dev), _dev(known, cuda_dev)
    dist, idx = ops.three_nn(ud, kd)
    np.testing.assert_array_equal(idx.cpu().numpy(), want_idx)
    np.testing.assert_array_equal(dist.cpu().numpy(), np.sqrt(want_d2))
    if refcuda.available():
        rd2, ridx = refcuda.three_nn(ud, kd)
        assert torch.equal(idx, ridx)
        assert torch.equal(dist, torch.sqrt(rd2))
    if m < 3:
        return
    # interpolation with the reference's inverse-distance weights (upsampling.py:97-100)
    C = 19
    g = torch.Generator().manual_seed(5)
    feat = torch.randn(B, C, m, generator=g)
    recip = 1.0 / (dist + 1e-8)
    w = (recip / recip.sum(dim=2, keepdim=True)).contiguous()
    fd = _dev(feat, cuda_dev).requires_grad_(True)
    out = ops.three_interpolate(fd, idx, w)
    want = orc.three_interp(feat.numpy(), want_idx, w.cpu().numpy())
    np.testing.assert_array_equal(out.detach().cpu().numpy(), want)
    go = torch.randn(B, C, n, generator=g)
    out.backward(_dev(go, cuda_dev))
    want_g = orc.three_interp_grad(go.numpy(), want_idx, w.cpu().numpy(), m)
    _assert_close_rel(fd.grad.cpu().numpy(), want_g)
    if refcuda.available():
        assert torch.equal(out.detach(), refcuda.three_interp(fd.detach(), idx, w))


@pytest.mark.parametrize("B,N,M,K,kind", [(2, 1024, 512, 32, "ball"), (2, 2048, 128, 24, "shell"), (1, 500, 77, 64, "quant"),
                                          (2, 300, 300, 16, "dup"), (1, 40, 10, 33, "ball"), (1, 5000, 100, 128, "shell")])
def test_knn_vs_oracle_and_torch(cuda_dev, B, N, M, K, kind):
    xyz = make_clouds(B, N, kind, seed=K)
    q = _queries(xyz, M)
    want_idx, want_d2 = orc.knn(K, xyz.numpy(), q.numpy())
    idx, d2 = ops.knn_query(K, _dev(xyz, cuda_dev), _dev(q, cuda_dev), return_dist2=True)
    np.testing.assert_array_equal(idx.cpu().numpy(), want_idx)
    np.testing.assert_array_equal(d2.cpu().numpy(), want_d2)
    if kind in ("ball", "shell") and K <= N:
        # parity with the reference's KNN (torch.cdist + topk, group.py:26-28) is "unpinned": compare the
        # neighbour SETS away from ties
        dist = torch.cdist(_dev(xyz, cuda_dev), _dev(q, cuda_dev))
        tk = dist.topk(k=K, dim=1, largest=False).indices.transpose(1, 2)
        a = np.sort(idx.cpu().numpy(), axis=-1)
        b = np.sort(tk.cpu().numpy(), axis=-1)
        assert (a == b).mean() > 0.999


def test_oracle_vs_compiled_reference_pins_the_oracle(cuda_dev):
    """Pins the CPU restatement against the reference's own kernels run here (the
    reference ships no golden vectors)."""
    if not refcuda.available():
        pytest.skip("oracle/_ref not built")
    for kind, N, m in (("quant", 1024, 256), ("dup", 1200, 300), ("quant", 300, 300), ("shell", 5000, 500), ("same", 64, 16)):
        xyz = make_clouds(2, N, kind, seed=11)
        xd = _dev(xyz, cuda_dev)
        np.testing.assert_array_equal(orc.fps(xyz.numpy(), m), refcuda.fps(xd, m).cpu().numpy())
        q = xyz[:, : N // 4].contiguous()
        np.testing.assert_array_equal(orc.ball_query(0.2, 16, xyz.numpy(), q.numpy()),
                                      refcuda.ball_query(0.2, 16, xd, _dev(q, cuda_dev)).cpu().numpy())
        d2, ix = orc.three_nn(xyz.numpy(), q.numpy())
        rd2, rix = refcuda.three_nn(xd, _dev(q, cuda_dev))
        np.testing.assert_array_equal(ix, rix.cpu().numpy())
        np.testing.assert_array_equal(d2, rd2.cpu().numpy())
        g = torch.Generator().manual_seed(N)
        feat = torch.randn(2, 9, N // 4, generator=g)
        w = torch.rand(2, N, 3, generator=g)
        np.testing.assert_array_equal(orc.three_interp(feat.numpy(), ix, w.numpy()),
                                      refcuda.three_interp(_dev(feat, cuda_dev), rix, _dev(w, cuda_dev)).cpu().numpy())


def test_cpu_tensor_raises():
    from adaptpoint_b200 import _cabi
    with pytest.raises(_cabi.ApbError):
        ops.furthest_point_sample(torch.zeros(1, 8, 3), 2)


@pytest.mark.parametrize("B,C,N,M,K", [(4, 32, 1024, 256, 32), (2, 7, 1200, 300, 24), (3, 64, 24000, 6000, 32),
                                       (1, 5, 100000, 2500, 32), (2, 3, 31, 31, 3), (2, 128, 120000, 1000, 8)])
def test_group_bf16_bit_exact(cuda_dev, B, C, N, M, K):
    """apb_group_bf16: a bit-exact copy of the bf16-rounded features, incl. odd sizes (scalar path), rows
    that fill shared memory (N = 100000) and rows longer than shared memory (direct gather)."""
    from adaptpoint_b200 import ops
    gen = torch.Generator().manual_seed(N + C)
    feat = torch.randn(B, C, N, generator=gen).to(torch.bfloat16).to(cuda_dev)
    idx = torch.randint(0, N, (B, M, K), generator=gen, dtype=torch.int32).to(cuda_dev)
    got = ops.grouping_operation_bf16(feat, idx)
    want = ops.torch_grouping_operation(feat.float(), idx).to(torch.bfloat16)
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))
    # and against the fp32 kernel on the same (bf16-representable) values
    got32 = ops.grouping_operation(feat.float().contiguous(), idx)
    assert torch.equal(got.float(), got32)


@pytest.mark.parametrize("B,N,M,radius,K,kind", [(4, 16384, 4096, 0.1, 32, "shell"), (2, 24000, 6000, 0.1, 32, "ball"),
                                                 (2, 6000, 6000, 0.2, 32, "ball"), (3, 5000, 1250, 0.05, 16, "quant"),
                                                 (2, 4096, 4096, 0.8, 24, "ball"), (2, 4100, 1000, 0.15, 64, "dup"),
                                                 (1, 100000, 25000, 0.1, 32, "shell"), (2, 4096, 512, 0.1, 32, "same"),
                                                 (2, 8192, 2048, 0.0, 8, "ball"), (2, 8192, 2048, 3.0, 96, "ball")])
def test_ball_query_grid_bit_exact(cuda_dev, B, N, M, radius, K, kind):
    """csrc/query_grid.cu (cell grid, nsample smallest in-radius indices) against the tiled index-order kernel and, at
    sizes the C oracle finishes quickly, against the oracle: identical rows on shells, balls, quantised coordinates
    (exact distance ties at the radius), duplicated points, an all-identical cloud, radius 0 and a radius that covers
    the whole cloud, queries that are / are not a subset of the support set."""
    from adaptpoint_b200 import _cabi, ops
    from adaptpoint_b200.synth import make_clouds
    gen = torch.Generator().manual_seed(N + K)
    if kind in ("shell", "ball"):
        xyz = make_clouds(B, N, kind, 5)
    elif kind == "quant":
        xyz = (make_clouds(B, N, "ball", 6) * 64).round() / 64
    elif kind == "dup":
        xyz = make_clouds(B, N, "ball", 7)
        xyz[:, N // 2:] = xyz[:, :N - N // 2]
    else:
        xyz = torch.full((B, N, 3), 0.25)
    if kind in ("ball", "quant") and M != N:
        q = torch.rand(B, M, 3, generator=gen) * 2.4 - 1.2          # queries outside the support's bounding box too
    else:
        sel = torch.stack([torch.randperm(N, generator=gen)[:M] for _ in range(B)])
        q = torch.gather(xyz, 1, sel.unsqueeze(-1).expand(-1, -1, 3)).contiguous()
    xd, qd = xyz.to(cuda_dev), q.to(cuda_dev)
    assert N >= ops.BALL_QUERY_GRID_MIN_N
    got = ops.ball_query(radius, K, xd, qd)
    want = torch.empty_like(got)
    _cabi.call("apb_ball_query_f32", [qd, xd, want, B, N, M, float(radius), K], cuda_dev)
    assert torch.equal(got, want)
    if N * M * B <= 3e8:
        from oracle import cpu as orc
        assert np.array_equal(got.cpu().numpy(), orc.ball_query(radius, K, xyz.numpy(), q.numpy()))
