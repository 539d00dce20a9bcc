"""Debug probe (GPU): which of eager / eager-serial / graph disagrees on the classifier loss on the tc backend?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import test_train_gpu as T
from adaptpoint_b200 import tcconv

dev = torch.device("cuda:0")
pts, y = T._batch(dev, 5)


def run(kind, conv="tc", **kw):
    tr = T._trainer(dev, conv, **kw)
    torch.manual_seed(1)
    script = [tr.draw_noise(pts.shape[0], pts.shape[1]) for _ in range(4)]
    if kind == "graph":
        T._scripted_noise(tr, script[0:3] + [script[3]])
        tr.capture_graph(pts, y, warmup=2)
    else:
        T._scripted_noise(tr, script[1:3] + [script[3]])
        for _ in range(2):
            tr.step(pts.clone(), y)
    torch.cuda.synchronize()
    return tr


base = run("eager")
for name, kind, kw in (("eager#2", "eager", {}), ("eager-serial", "eager", {"overlap": False}), ("graph", "graph", {}),
                       ("graph-serial", "graph", {"overlap": False})):
    tr = run(kind, **kw)
    T._copy_state(base, tr)
    # `base` must not move: run the candidate first on a copy of base's state, then compare with a fresh copy
    out = torch.stack([o.float() for o in tr.step(pts.clone(), y)]).cpu()
    print(name, out.tolist(), flush=True)
ref = run("eager")
T._copy_state(base, ref)
print("eager(ref)", torch.stack([o.float() for o in ref.step(pts.clone(), y)]).cpu().tolist())
if os.environ.get("NOSTATS"):
    pass
