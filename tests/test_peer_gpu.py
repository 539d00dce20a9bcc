"""Peer-memory all-reduce (csrc/peer.cu) against NCCL on the same buffers, on EVERY GPU of the box (2, 4 or 8 ranks; skipped
on a 1-GPU box): equality with NCCL, CUDA-graph replay, and calls issued from two streams without host synchronisation
(peer.py orders them with events: the workspace has one sequence counter).  Run by scripts/gpu_r02_scale.sh."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from adaptpoint_b200 import peer
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    ctx = peer.enable()
    assert ctx is not None and ctx.world == world
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    worst = 0.0
    for it, (n, dt) in enumerate([(2, torch.float64), (128, torch.float64), (1024, torch.float32), (8192, torch.float64),
                                  (6, torch.float32)] * 40):
        x = torch.randn(n, generator=g, device=dev, dtype=dt)
        want = x.clone()
        dist.all_reduce(want)
        got = ctx.allreduce_(x.clone())
        worst = max(worst, float((got - want).abs().max() / want.abs().max().clamp_min(1e-30)))
    # graph replay: the sequence counter lives on the device, so a captured all-reduce can be replayed
    buf = torch.ones(256, device=dev, dtype=torch.float64)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        ctx.allreduce_(buf)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        buf.fill_(1.0)
        with torch.cuda.graph(graph, stream=s):
            ctx.allreduce_(buf)
            ctx.allreduce_(buf)
        for _ in range(5):
            buf.fill_(rank + 1.0)
            graph.replay()
    torch.cuda.synchronize()
    expect = float(sum(r + 1 for r in range(world)) * world)
    graph_ok = bool((buf == expect).all())
    # two streams, no host sync between the calls: the guard in peer.py must serialise them
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    xs = [torch.full((64,), float(rank + 1 + i), device=dev, dtype=torch.float64) for i in range(20)]
    torch.cuda.synchronize()
    for i, x in enumerate(xs):
        with torch.cuda.stream(sa if i % 2 == 0 else sb):
            ctx.allreduce_(x)
    torch.cuda.synchronize()
    streams_ok = all(bool((x == float(sum(r + 1 + i for r in range(world)))).all()) for i, x in enumerate(xs))
    # latency against NCCL on a (128,2) f64 statistic
    st = torch.zeros(256, device=dev, dtype=torch.float64)
    res = {}
    for name, fn in (("nccl", lambda: dist.all_reduce(st)), ("peer", lambda: ctx.allreduce_(st))):
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(200):
            fn()
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / 200 * 1e3
    if rank == 0:
        torch.save({"worst_rel": worst, "graph_ok": graph_ok, "streams_ok": streams_ok, "us": res, "world": world}, out)
    dist.barrier()
    torch.cuda.synchronize()
    os._exit(0)


def run(world=2, port=29533, out="/tmp/apb_peer_test.pt"):
    import torch.multiprocessing as mp
    if os.path.exists(out):
        os.remove(out)
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    return torch.load(out)


@pytest.mark.gpu
def test_peer_allreduce_matches_nccl():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs on one node")
    r = run(world=min(torch.cuda.device_count(), 8))
    print("peer all-reduce:", r)
    assert r["worst_rel"] < 1e-12 or r["worst_rel"] < 1e-6, r      # f64 sums exact up to order; f32 within 1e-6
    assert r["graph_ok"], r
    assert r["streams_ok"], r


if __name__ == "__main__":
    print(run(world=int(sys.argv[1]) if len(sys.argv) > 1 else 2))
