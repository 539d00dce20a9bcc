"""Where does run-to-run variation of the classifier's training loss come from?  Two identical serial runs."""
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


def run():
    torch.manual_seed(0)
    torch.cuda.manual_seed(0)
    tr = AdaptPointTrainer(device=dev, seed=0, conv_backend="torch", overlap=False)
    torch.manual_seed(1)
    torch.cuda.manual_seed(1)
    out = {}
    noise = tr.draw_noise(8, 2048)
    fake, gl, dl = tr.gan_iteration(pts.clone(), y, noise)
    out["fake1"] = fake.clone()
    out["G1"] = torch.cat([p.detach().flatten() for p in tr.generator.parameters()])
    torch.cuda.manual_seed(7)
    logits, loss = tr.classifier_iteration(fake, y, noise["keep"])
    out["logits1"], out["loss1"] = logits.clone(), loss.clone()
    out["C1"] = torch.cat([p.detach().flatten() for p in tr.classifier.parameters()])
    out["Cbuf1"] = torch.cat([b.detach().flatten().float() for b in tr.classifier.buffers()])
    tr.classifier.eval()
    with torch.no_grad():
        data = {"pos": pts[:, :1024, :3].contiguous(), "x": pts[:, :1024, :4].transpose(1, 2).contiguous()}
        out["eval_logits"] = tr.classifier(data).clone()
    tr.classifier.train()
    torch.cuda.manual_seed(9)
    with torch.no_grad():
        out["train_logits_fixed_seed"] = tr.classifier(data).clone()
    return out


a, b = run(), run()
for k in a:
    d = (a[k].double() - b[k].double()).abs()
    print(f"{k:26s} max|diff| {float(d.max()):.3e}   max|a| {float(a[k].abs().max()):.3e}   n_diff>1e-4 {int((d > 1e-4).sum())}")
