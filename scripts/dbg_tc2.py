"""Is the tc-vs-fp32 gradient gap numerics (bf16 operand rounding) or a kernel bug?  Emulate the kernels' numerics in
torch (operands rounded to bf16, fp32 math) and compare all three."""
import sys, torch
sys.path.insert(0, '/root/repo')
from adaptpoint_b200 import tcconv
from adaptpoint_b200.models import build_model
from adaptpoint_b200.synth import make_clouds
import torch.nn.functional as F
dev = torch.device('cuda:0')
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False

class Emu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        ctx.save_for_backward(x, w); ctx.hb = b is not None
        B, cin = x.shape[:2]; cout = w.shape[0]
        xr = x.reshape(B, cin, -1).bfloat16().float(); wr = w.reshape(cout, cin).bfloat16().float()
        y = torch.einsum('oi,bip->bop', wr.double(), xr.double()).float().contiguous()
        if b is not None: y = y + b.view(1, -1, 1)
        return y.view((B, cout) + x.shape[2:])
    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        B, cin = x.shape[:2]; cout = w.shape[0]
        g = gy.reshape(B, cout, -1)
        gx = torch.einsum('oi,bop->bip', w.reshape(cout, cin).bfloat16().double(), g.bfloat16().double()).float().contiguous().view(x.shape)
        gw = torch.einsum('bop,bip->oi', g.double(), x.reshape(B, cin, -1).double()).float().view(w.shape)
        return gx, gw, (g.sum((0, 2)) if ctx.hb else None)

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
torch.manual_seed(0)
model = build_model("pointnext-s").to(dev).train()
xyz = make_clouds(B, 1024, "shell", 3).to(dev)
x = torch.cat([xyz, xyz[:, :, 1:2] - xyz[:, :, 1:2].amin(1, keepdim=True)], -1).transpose(1, 2).contiguous()
y = torch.arange(B, device=dev) % 15
def run(mode):
    model.zero_grad(); torch.manual_seed(7)
    if mode == 'emu':
        orig = tcconv.PointwiseConvFn.apply
        tcconv.PointwiseConvFn.apply = Emu.apply
        try:
            with tcconv.conv_backend('tc'):
                logits, loss = model.get_logits_loss({"pos": xyz, "x": x}, y); loss.backward()
        finally:
            tcconv.PointwiseConvFn.apply = orig
    else:
        with tcconv.conv_backend(mode):
            logits, loss = model.get_logits_loss({"pos": xyz, "x": x}, y); loss.backward()
    return {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None}
g = {m: run(m) for m in ('torch', 'tc', 'emu')}
cos = lambda a, b: round(F.cosine_similarity(a.flatten(), b.flatten(), dim=0).item(), 4)
for n in g['torch']:
    if g['torch'][n].numel() >= 2048 and 'convs' in n:
        print(n, 'tc~fp32', cos(g['tc'][n], g['torch'][n]), 'emu~fp32', cos(g['emu'][n], g['torch'][n]), 'tc~emu', cos(g['tc'][n], g['emu'][n]))
