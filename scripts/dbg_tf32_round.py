"""Probe (GPU): does the TMA-fed TF32 path truncate or round its fp32 operands?  x = 1 + 3*2^-12 is not a tf32 number:
round-to-nearest gives 1 + 2^-10, truncation gives 1."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from adaptpoint_b200 import _cabi, tcconv
dev = torch.device("cuda:0")
B, cin, cout, P = 1, 64, 32, 256
x = torch.full((B, cin, P), 1 + 3 * 2.0 ** -12, device=dev)
w = torch.ones(cout, cin, device=dev)
wb, kp = tcconv.pack_f32(w)
y = torch.empty(B, cout, P, device=dev)
for entry in ("apb_pwconv_fwd_tf32", "apb_pwconv_fwd_tf32_rows"):
    _cabi.call(entry, [x, wb, None, None, 0, y, B, cin, cout, P, kp], dev)
    print(entry, "x operand:", y[0, 0, 0].item() / cin, "(1.0 = truncated, 1.0009765625 = rounded to nearest)")
w2 = torch.full((cout, cin), 1 + 3 * 2.0 ** -12, device=dev)
wb2, _ = tcconv.pack_f32(w2)
_cabi.call("apb_pwconv_fwd_tf32", [torch.ones_like(x), wb2, None, None, 0, y, B, cin, cout, P, kp], dev)
print("w operand:", y[0, 0, 0].item() / cin)
gw = torch.zeros(cout, cin, device=dev)
_cabi.call("apb_pwconv_wgrad_tf32", [torch.ones(B, cout, P, device=dev), x, gw, B, cin, cout, P], dev)
print("wgrad x operand:", gw[0, 0].item() / P)
