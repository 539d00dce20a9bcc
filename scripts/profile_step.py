"""Where does one AdaptPoint step go?  Phase timings (CUDA events + host wall) and a torch.profiler kernel table.
Run on the GPU box:  python scripts/profile_step.py > gpurun_out/step_profile.txt"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import loader_batch, B_PER_GPU, N_POINTS  # noqa: E402
from adaptpoint_b200.train import AdaptPointTrainer  # noqa: E402
from adaptpoint_b200 import tcconv  # noqa: E402

dev = torch.device("cuda:0")
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
tr = AdaptPointTrainer(device=dev, seed=0)
pts, y = loader_batch(B_PER_GPU, N_POINTS, 100)
pts, y = pts.to(dev), y.to(dev)
for _ in range(3):
    tr.step(pts.clone(), y)
torch.cuda.synchronize()
tcconv.set_conv_backend(tr.conv_backend)   # the direct phase calls below bypass tr.step's context


def timed(fn, n=5):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        out = fn()
    e1.record()
    host = (time.perf_counter() - t0) / n * 1e3
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, host, out


g_ms, g_host, (fake, _, _) = timed(lambda: tr.gan_iteration(pts.clone(), y))
c_ms, c_host, _ = timed(lambda: tr.classifier_iteration(fake, y))
print(f"gan_iteration        gpu {g_ms:8.2f} ms   host-enqueue {g_host:8.2f} ms")
print(f"classifier_iteration gpu {c_ms:8.2f} ms   host-enqueue {c_host:8.2f} ms")

# finer: generator forward, discriminator forward, classifier eval forward, classifier train fwd / bwd
G, D, C = tr.generator, tr.discriminator, tr.classifier
inp = pts[:, :, :3].contiguous()
G.train(); D.train()
ms, host, (_, gen) = timed(lambda: G(inp))
print(f"  G forward            gpu {ms:8.2f}  host {host:8.2f}")
ms, host, _ = timed(lambda: D(gen.detach()))
print(f"  D forward            gpu {ms:8.2f}  host {host:8.2f}")


def g_fb():
    _, gen = G(inp)
    loss = D(gen).mean()
    loss.backward()
    G.zero_grad(); D.zero_grad()


ms, host, _ = timed(g_fb)
print(f"  G fwd + D fwd + bwd  gpu {ms:8.2f}  host {host:8.2f}")
data = {"pos": inp[:, :1024].contiguous(), "x": pts[:, :1024, :4].transpose(1, 2).contiguous()}
C.eval()
with torch.no_grad():
    ms, host, _ = timed(lambda: C(data))
print(f"  C eval forward N=1024 (no grad) gpu {ms:8.2f}  host {host:8.2f}")
data2 = {"pos": inp, "x": pts[:, :, :4].transpose(1, 2).contiguous()}
with torch.no_grad():
    ms, host, _ = timed(lambda: C(data2))
print(f"  C eval forward N=2048 (no grad) gpu {ms:8.2f}  host {host:8.2f}")
C.train()


def c_fb():
    logits, loss = C.get_logits_loss(data, y)
    loss.backward()
    C.zero_grad()


ms, host, _ = timed(c_fb)
print(f"  C train fwd+bwd N=1024 gpu {ms:8.2f}  host {host:8.2f}")

from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
    tr.step(pts.clone(), y)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=70))
print(prof.key_averages(group_by_input_shape=True).table(sort_by="cuda_time_total", row_limit=60,
                                                          max_name_column_width=50, max_shapes_column_width=90))
