"""Debug probe (GPU): where does the discriminator's input gradient vanish on the tc backend?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import _model_cases as mc
from adaptpoint_b200 import tcconv
from adaptpoint_b200.imitator import PointDiscriminator1
from fill import fill_state

dev = torch.device("cuda:0")
mc.strict_fp32()
for backend in ("torch", "tc"):
    dis = PointDiscriminator1(num_classes=15, normal_channel=False)
    fill_state(dis, 3)
    dis.to(dev).eval()
    pos, _, _ = mc.batch(2, 512, 14)
    pos_g = pos.to(dev).requires_grad_(True)
    acts = {}
    def hook(name):
        def f(mod, inp, out):
            if isinstance(out, torch.Tensor) and out.requires_grad:
                out.register_hook(lambda g, n=name: acts.__setitem__(n, float(g.norm())))
        return f
    for n, m in dis.named_modules():
        if n:
            m.register_forward_hook(hook(n))
    with tcconv.conv_backend(backend):
        prob = dis(pos_g)
        prob.sum().backward()
    print(backend, "prob", prob.flatten().tolist(), "grad_pos norm", float(pos_g.grad.norm()))
    print({k: round(v, 6) for k, v in acts.items()})
