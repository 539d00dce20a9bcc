"""Data-parallel path on CPU: world_size 2, gloo.  Checks the explicit gradient all-reduce of train.py
(allreduce_grads) and that two ranks stepping AdaptPointTrainer on different shards end with identical parameters
(same initial broadcast, averaged gradients, SyncBatchNorm) -- the N>1 host logic of bench.py --gpus N."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from adaptpoint_b200 import train
        # (1) allreduce_grads averages
        p = [torch.nn.Parameter(torch.zeros(3)), torch.nn.Parameter(torch.zeros(2, 2))]
        p[0].grad = torch.full((3,), float(rank + 1))
        p[1].grad = torch.full((2, 2), float(10 * (rank + 1)))
        train.allreduce_grads(p)
        assert torch.allclose(p[0].grad, torch.full((3,), 1.5)) and torch.allclose(p[1].grad, torch.full((2, 2), 15.0))
        # (2) two ranks, different shards, one step -> identical parameters
        from oracle import torch_ops
        from adaptpoint_b200.synth import make_clouds
        with torch_ops.patched():
            tr = train.AdaptPointTrainer(device="cpu", seed=rank)     # different seeds: the broadcast must align them
            B, N = 2, 256
            pos = make_clouds(B, N, "shell", 10 + rank)
            pts = torch.cat([pos, pos[:, :, 1:2] - pos[:, :, 1:2].amin(1, keepdim=True)], -1)
            y = torch.tensor([rank, rank + 3])
            losses = tr.step(pts, y)
            assert all(torch.isfinite(l) for l in losses)
        flat = torch.cat([q.detach().flatten() for net in (tr.classifier, tr.generator, tr.discriminator)
                          for q in net.parameters()])
        torch.save(flat, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_world_size_2_gloo(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a = torch.load(tmp_path / "rank0.pt")
    b = torch.load(tmp_path / "rank1.pt")
    assert a.shape == b.shape and torch.isfinite(a).all()
    assert torch.allclose(a, b, rtol=1e-5, atol=1e-6), (a - b).abs().max()
