"""Natural run-to-run spread of the step losses (fp32 atomics) vs the serial/overlap difference."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from adaptpoint_b200.synth import make_clouds  # noqa: E402
from adaptpoint_b200.train import AdaptPointTrainer  # noqa: E402

dev = torch.device("cuda:0")
pos = make_clouds(8, 2048, "shell", 5)
pts = torch.cat([pos, pos[:, :, 1:2] - pos[:, :, 1:2].amin(1, keepdim=True)], -1).to(dev)
y = (torch.arange(8) % 15).to(dev)
for overlap in (False, False, True, True, False, True):
    torch.manual_seed(0)
    torch.cuda.manual_seed(0)
    tr = AdaptPointTrainer(device=dev, seed=0, conv_backend="torch", overlap=overlap)
    torch.manual_seed(1)
    torch.cuda.manual_seed(1)
    rows = []
    for _ in range(3):
        rows.append([round(float(o), 5) for o in tr.step(pts.clone(), y)])
    print("overlap" if overlap else "serial ", rows, flush=True)
