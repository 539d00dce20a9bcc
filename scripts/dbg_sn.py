import os, sys, copy
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.nn as nn
from adaptpoint_b200 import spectral, tcconv
dev = torch.device("cuda:0")
rel = lambda a, b: ((a.double().cpu() - b.double().cpu()).abs().max() / b.double().cpu().abs().max()).item()
for name, make in (("fc1", lambda: nn.Linear(1024, 512)), ("conv1", lambda: nn.Conv2d(3, 64, 1))):
    torch.manual_seed(3)
    base = make()
    ref = nn.utils.parametrizations.spectral_norm(copy.deepcopy(base)).to(dev).train()
    ours = spectral.spectral_norm(copy.deepcopy(base)).to(dev).train()
    ours.load_state_dict(ref.state_dict())
    with tcconv.conv_backend("tc"):
        for step in range(3):
            sd = {k: v.detach().cpu().double() for k, v in ref.state_dict().items()}
            sdo = {k: v.detach().cpu().double() for k, v in ours.state_dict().items()}
            W = sd['parametrizations.weight.original'].flatten(1); v0 = sd['parametrizations.weight.0._v']; vo = sdo['parametrizations.weight.0._v']
            t = W @ v0; u = t / t.norm(); s = W.t() @ u; sigma = s.norm(); want = W / sigma
            w_ref, w_ours = ref.weight, ours.weight
            print(name, step, "state v dev", (v0 - vo).abs().max().item(), "torch vs f64", rel(w_ref.flatten(1), want), "ours vs f64", rel(w_ours.flatten(1), want),
                  "ours vs torch", rel(w_ours, w_ref), flush=True)
