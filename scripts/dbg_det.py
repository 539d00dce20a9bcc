"""Debug probe (GPU): run-to-run spread of the classifier's training-mode forward on the tc backend, B=8."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from adaptpoint_b200 import tcconv, norm
from adaptpoint_b200.models import build_model
from adaptpoint_b200.synth import make_clouds
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = build_model("pointnext-s").to(dev).train()
for m in model.modules():
    if isinstance(m, torch.nn.Dropout):
        m.p = 0.0
xyz = make_clouds(8, 1024, "shell", 3).to(dev)
x = torch.cat([xyz, xyz[:, :, 1:2] - xyz[:, :, 1:2].amin(1, keepdim=True)], -1).transpose(1, 2).contiguous()
y = torch.arange(8, device=dev) % 15
state = {k: v.clone() for k, v in model.state_dict().items()}
for tag, fuse in (("fused-stats", True), ("separate-stats", False)):
    orig = tcconv.conv_with_stats
    if not fuse:
        tcconv.conv_with_stats = lambda mod, t: (mod(t), None)
    vals, feats = [], []
    for i in range(4):
        model.load_state_dict(state)
        with tcconv.conv_backend("tc"):
            f = model.encoder.forward_cls_feat({"pos": xyz, "x": x})
            logits, loss = model.get_logits_loss({"pos": xyz, "x": x}, y)
        vals.append(loss.item()); feats.append(f.detach().clone())
    tcconv.conv_with_stats = orig
    spread = max((feats[0] - g).abs().max().item() for g in feats[1:]) / feats[0].abs().max().item()
    print(tag, "loss", vals, "encoder feature rel spread", spread, flush=True)
