"""Debug probe (GPU): per-parameter generator gradient norms, strict (torch backend) vs tc, on the step golden's inputs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import _model_cases as mc

dev = torch.device("cuda:0")
mc.strict_fp32()
res = {}
for backend in ("torch", "tc"):
    grads = {}
    orig = mc.check
    mc.begin("record")
    from adaptpoint_b200.train import AdaptPointTrainer
    real_probe = None
    import adaptpoint_b200.train as T
    old_probe = T.AdaptPointTrainer._probe
    def probe(self, name, module, grads=grads):
        if name == "generator":
            for n, p in module.named_parameters():
                if p.grad is not None:
                    grads[n] = p.grad.detach().double().cpu().clone()
        return old_probe(self, name, module)
    T.AdaptPointTrainer._probe = probe
    try:
        mc.case_step(dev, backend)
    finally:
        T.AdaptPointTrainer._probe = old_probe
    res[backend] = grads
a, b = res["torch"], res["tc"]
rows = []
for n in a:
    na, nb = a[n].norm().item(), b[n].norm().item()
    cos = float((a[n].flatten() @ b[n].flatten()) / max(na * nb, 1e-30))
    rows.append((abs(nb * nb - na * na), n, na, nb, cos))
rows.sort(reverse=True)
tot_a = sum(r[2] ** 2 for r in rows) ** 0.5; tot_b = sum(r[3] ** 2 for r in rows) ** 0.5
print("total", tot_a, tot_b)
for r in rows[:14]:
    print(f"{r[1]:70s} strict {r[2]:.5f} tc {r[3]:.5f} cos {r[4]:.4f}")
