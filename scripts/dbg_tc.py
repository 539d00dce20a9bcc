import torch, sys
sys.path.insert(0, '/root/repo')
from adaptpoint_b200 import tcconv
from adaptpoint_b200.models import build_model
from adaptpoint_b200.synth import make_clouds
dev = torch.device('cuda:0')
torch.manual_seed(0)
model = build_model("pointnext-s").to(dev).train()
xyz = make_clouds(8, 1024, "shell", 3).to(dev)
x = torch.cat([xyz, xyz[:, :, 1:2] - xyz[:, :, 1:2].amin(1, keepdim=True)], -1).transpose(1, 2).contiguous()
y = torch.arange(8, device=dev) % 15
feats = {}
def hook(name):
    def f(m, i, o):
        t = o[1] if isinstance(o, (list, tuple)) else o
        feats.setdefault(name, []).append(t.detach().clone())
    return f
for i, st in enumerate(model.encoder.encoder):
    st.register_forward_hook(hook(f"stage{i}"))
res = {}
for backend in ("torch", "tc"):
    torch.manual_seed(7)
    with tcconv.conv_backend(backend):
        logits, loss = model.get_logits_loss({"pos": xyz, "x": x}, y)
    res[backend] = (logits.detach(), loss.item())
for k, v in feats.items():
    a, b = v
    print(k, a.shape, 'rel', ((a-b).abs().max()/a.abs().max()).item(), 'cos', torch.nn.functional.cosine_similarity(a.flatten(), b.flatten(), dim=0).item())
a, b = res['torch'][0], res['tc'][0]
print('logits rel', ((a-b).abs().max()/a.abs().max()).item(), 'cos', torch.nn.functional.cosine_similarity(a.flatten(), b.flatten(), dim=0).item(), res['torch'][1], res['tc'][1])
grads = {}
for backend in ("torch", "tc"):
    model.zero_grad()
    torch.manual_seed(7)
    with tcconv.conv_backend(backend):
        logits, loss = model.get_logits_loss({"pos": xyz, "x": x}, y)
        loss.backward()
    grads[backend] = {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None}
for n, g in grads["torch"].items():
    gt = grads["tc"][n]
    if g.numel() < 64: continue
    print(n, tuple(g.shape), 'cos', round(torch.nn.functional.cosine_similarity(g.flatten(), gt.flatten(), dim=0).item(), 4), 'norm ratio', round((gt.norm()/g.norm()).item(), 3))
# same comparison with TF32 cuDNN/cuBLAS as a second opinion on how noisy this gradient is
torch.backends.cudnn.allow_tf32 = True; torch.backends.cuda.matmul.allow_tf32 = True
model.zero_grad(); torch.manual_seed(7)
logits, loss = model.get_logits_loss({"pos": xyz, "x": x}, y); loss.backward()
for n, p in model.named_parameters():
    if p.grad is not None and p.grad.numel() >= 64 and ('encoder.0.' in n or 'encoder.1.' in n):
        g = grads["torch"][n]
        print('TF32-cudnn', n, 'cos', round(torch.nn.functional.cosine_similarity(g.flatten(), p.grad.flatten(), dim=0).item(), 4))
