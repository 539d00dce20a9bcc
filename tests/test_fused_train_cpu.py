"""Host-side algebra of the fused two-conv set abstraction (adaptpoint_b200/fused_train.py: SaTwoLayerFn) on CPU: every
C-ABI call is swapped for its torch restatement (tests/_sa2_emul.py), and forward value + every gradient are compared
with plain torch autograd of the reference's op chain (pointnext.py:164-166, conv.py:24-60).  The CUDA kernels themselves
are compared with the same restatements in tests/test_fused_train_gpu.py."""
import pytest
import torch

from adaptpoint_b200 import _cabi, fused_train
import _sa2_emul as emul


def _case(seed, B=2, N=48, M=12, K=32, cin=8, C1=16, C2=24):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    p = r(B, N, 3) * 0.3
    q = p[:, :M].clone()
    f = r(B, cin, N)
    idx = torch.randint(0, N, (B, M, K), generator=g, dtype=torch.int32)
    w1 = r(C1, cin + 3, 1, 1) * 0.4
    w2 = r(C2, C1, 1, 1) * 0.4
    bn = lambda c: (1 + 0.3 * r(c), 0.2 * r(c), 0.1 * r(c), 0.5 + torch.rand(c, generator=g))
    return p, q, f, idx, w1, bn(C1), w2, bn(C2)


@pytest.mark.parametrize("training,relu2,coords", [(True, True, False), (True, False, False), (False, True, True),
                                                    (False, False, True)])
def test_two_conv_algebra_matches_autograd(monkeypatch, training, relu2, coords):
    monkeypatch.setattr(_cabi, "call", emul.call)
    monkeypatch.setattr(fused_train, "EXACT_GEMM", True)
    p, q, f, idx, w1, (g1, b1, rm1, rv1), w2, (g2, b2, rm2, rv2) = _case(3)
    # negative BatchNorm-2 weights exercise the min branch of the monotone max
    g2 = g2 * torch.where(torch.arange(g2.numel()) % 3 == 0, -1.0, 1.0)
    inv_r = 1.0 / 0.35
    leaves = [t.clone().requires_grad_(True) for t in (f, w1, g1, b1, w2, g2, b2)]
    pp, qq = p.clone().requires_grad_(coords), q.clone().requires_grad_(coords)
    rms = [t.clone() for t in (rm1, rv1, rm2, rv2)]
    out = fused_train.SaTwoLayerFn.apply(pp, qq, leaves[0], idx, leaves[1], leaves[2], leaves[3], rms[0], rms[1], 0.1, 1e-5,
                                         leaves[4], leaves[5], leaves[6], rms[2], rms[3], 0.1, 1e-5, relu2, training, None,
                                         inv_r)
    go = torch.randn(out.shape, generator=torch.Generator().manual_seed(9))
    out.backward(go)

    ref_leaves = [t.double().clone().requires_grad_(True) for t in (f, w1, g1, b1, w2, g2, b2)]
    rp, rq = p.double().clone().requires_grad_(coords), q.double().clone().requires_grad_(coords)
    rrm = [t.double().clone() for t in (rm1, rv1, rm2, rv2)]
    ref = emul.reference_block(rp, rq, ref_leaves[0], idx, ref_leaves[1], ref_leaves[2], ref_leaves[3], ref_leaves[4],
                               ref_leaves[5], ref_leaves[6], relu2, training, inv_r, rrm[0], rrm[1], rrm[2], rrm[3])
    ref.backward(go.double())

    def close(name, a, b, tol=2e-4):
        a, b = a.double(), b.double()
        err = (a - b).abs().max().item()
        assert err <= tol * (1.0 + b.abs().max().item()), f"{name}: max |diff| {err:.3e} vs scale {b.abs().max().item():.3e}"

    close("out", out, ref)
    for name, a, b in zip(("f", "w1", "g1", "b1", "w2", "g2", "b2"), leaves, ref_leaves):
        assert a.grad is not None, name
        close("grad " + name, a.grad, b.grad)
    if coords:
        close("grad p", pp.grad, rp.grad)
        close("grad q", qq.grad, rq.grad)
    if training:   # running statistics updated like torch's
        for name, a, b in zip(("rm1", "rv1", "rm2", "rv2"), rms, rrm):
            close(name, a, b)


# ---- SyncBatchNorm semantics of the fused two-conv path: two gloo ranks, each half of the batch, against one process on
# ---- the whole batch (the all-reduces of SaTwoLayerFn: both forward statistics, both backward sum pairs) --------------
def _sync_worker(rank, world, port, out_dir):
    import os
    import sys
    import torch.distributed as dist
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    sys.path.insert(0, os.path.join(root, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import _sa2_emul as em
        from adaptpoint_b200 import _cabi as cabi, fused_train as ft
        cabi.call = em.call
        ft.EXACT_GEMM = True
        p, q, f, idx, w1, (g1, b1, rm1, rv1), w2, (g2, b2, rm2, rv2) = _case(5, B=4)
        sl = slice(2 * rank, 2 * rank + 2)
        leaves = [t.clone().requires_grad_(True) for t in (f[sl], w1, g1, b1, w2, g2, b2)]
        rms = [t.clone() for t in (rm1, rv1, rm2, rv2)]
        out = ft.SaTwoLayerFn.apply(p[sl], q[sl], leaves[0], idx[sl], leaves[1], leaves[2], leaves[3], rms[0], rms[1], 0.1, 1e-5,
                                    leaves[4], leaves[5], leaves[6], rms[2], rms[3], 0.1, 1e-5, True, True, dist.group.WORLD,
                                    1.0 / 0.35)
        go = torch.randn((4,) + tuple(out.shape[1:]), generator=torch.Generator().manual_seed(9))
        out.backward(go[sl])
        torch.save({"out": out.detach(), "grads": [t.grad for t in leaves], "rms": rms},
                   os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_conv_syncbn_two_ranks_equal_one_process(monkeypatch, tmp_path):
    import socket
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_sync_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = torch.load(tmp_path / "rank0.pt"), torch.load(tmp_path / "rank1.pt")

    monkeypatch.setattr(_cabi, "call", emul.call)
    monkeypatch.setattr(fused_train, "EXACT_GEMM", True)
    p, q, f, idx, w1, (g1, b1, rm1, rv1), w2, (g2, b2, rm2, rv2) = _case(5, B=4)
    leaves = [t.clone().requires_grad_(True) for t in (f, w1, g1, b1, w2, g2, b2)]
    rms = [t.clone() for t in (rm1, rv1, rm2, rv2)]
    out = fused_train.SaTwoLayerFn.apply(p, q, leaves[0], idx, leaves[1], leaves[2], leaves[3], rms[0], rms[1], 0.1, 1e-5,
                                         leaves[4], leaves[5], leaves[6], rms[2], rms[3], 0.1, 1e-5, True, True, None,
                                         1.0 / 0.35)
    go = torch.randn(out.shape, generator=torch.Generator().manual_seed(9))
    out.backward(go)

    def close(name, a, b):
        err = (a.double() - b.double()).abs().max().item()
        assert err <= 2e-4 * (1.0 + b.double().abs().max().item()), f"{name}: {err:.3e}"

    close("out", torch.cat([r0["out"], r1["out"]]), out)
    close("grad f", torch.cat([r0["grads"][0], r1["grads"][0]]), leaves[0].grad)
    for i, name in enumerate(("w1", "g1", "b1", "w2", "g2", "b2"), start=1):
        # parameter gradients are per-rank sums: the ranks' sum is the whole-batch gradient
        close("grad " + name, r0["grads"][i] + r1["grads"][i], leaves[i].grad)
    for i, name in enumerate(("rm1", "rv1", "rm2", "rv2")):
        close(name, r0["rms"][i], rms[i])
        close(name + " (rank 1)", r1["rms"][i], rms[i])
