"""wgrad microbenchmark over the step's shapes: sweep the 32-point boxes per pipeline stage (mt field = 10 + kb).  JSON lines."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from adaptpoint_b200 import _cabi
from bench import ClockSampler, peaks
dev = torch.device("cuda:0")
hbm, _ = peaks()
SHAPES = [(32, 256, 512, 64), (32, 64, 128, 2048), (32, 3, 64, 2048), (32, 256, 512, 2048), (32, 128, 256, 4096),
          (32, 131, 128, 4096), (32, 1536, 512, 256), (32, 64, 128, 8192), (32, 32, 64, 16384), (32, 768, 256, 512),
          (32, 259, 256, 2048), (32, 35, 32, 16384), (32, 384, 128, 1024), (32, 256, 512, 512), (32, 128, 512, 2048),
          (8, 64, 256, 24000), (8, 256, 64, 24000), (8, 512, 512, 6000)]
flush = torch.empty(256 * 1024 * 1024 // 4, device=dev)
lib = _cabi.load()
sampler = ClockSampler(0); sampler.start()
rows = []
for B, cin, cout, P in SHAPES:
    x = torch.randn(B, cin, P, device=dev); gy = torch.randn(B, cout, P, device=dev)
    gw = torch.zeros(cout, cin, device=dev)
    want = torch.einsum("bop,bip->oi", gy.double(), x.double()).float()
    for mt in (10, 11, 12, 14):           # 10 + boxes per stage (10 = automatic); MT stays automatic (= 1)
        for mc in (16,):
            if lib.apb_tune_wgrad(mt, mc) != 0 or lib.apb_tune_wgrad(0, mc) != 0:
                continue
            try:
                gw.zero_(); _cabi.call("apb_pwconv_wgrad_tf32", [gy, x, gw, B, cin, cout, P], dev)
            except _cabi.ApbError:
                continue
            err = ((gw - want).abs().max() / want.abs().max()).item()
            ts = []
            for _ in range(8):
                flush.add_(1.0); gw.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); _cabi.call("apb_pwconv_wgrad_tf32", [gy, x, gw, B, cin, cout, P], dev); b.record()
                torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
            us = sorted(ts)[len(ts) // 2] * 1e3
            rows.append({"shape": [B, cin, cout, P], "mt": mt, "min_chunks": mc, "us": round(us, 2), "rel_err": round(err, 6),
                         "hbm_frac": round(4.0 * B * P * (cin + cout) / us / 1e3 / hbm, 3)})
lib.apb_tune_wgrad(10, 0)
lib.apb_tune_wgrad(0, 0)
clocks = sampler.stop()
for r in rows:
    r["clocks"] = clocks
    print(json.dumps(r))
