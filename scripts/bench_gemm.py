"""Microbenchmark of the 1x1-conv GEMM kernels (TF32): TMA-fed tc_cm_gemm vs the round-1 register-staged rows kernel,
on the step's and the XL model's shapes.  One JSON line per (shape, kernel): us, achieved GB/s by the algorithmic bytes
4*B*P*(Cin+Cout), fraction of the measured HBM peak, TFLOP/s; clocks sampled during the run."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from adaptpoint_b200 import _cabi, tcconv
from bench import ClockSampler, peaks

SHAPES = [(32, 128, 512, 2048), (32, 512, 128, 2048), (32, 35, 32, 16384), (32, 32, 64, 16384), (32, 67, 64, 8192),
          (32, 64, 128, 8192), (32, 131, 128, 4096), (32, 128, 256, 4096), (32, 259, 256, 2048), (32, 256, 512, 2048),
          (32, 64, 64, 2048), (32, 3, 64, 2048), (32, 128, 1024, 2048),
          (8, 512, 512, 65536), (8, 64, 256, 24000), (8, 256, 64, 24000), (8, 128, 512, 6000), (8, 512, 128, 6000),
          (8, 1024, 256, 1500)]


def timeit(fn, flush, iters=20):
    for _ in range(3):
        fn()
    evs = []
    for _ in range(iters):
        flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    ts = sorted(x.elapsed_time(y) for x, y in evs)
    return ts[len(ts) // 2] * 1e3


def main():
    dev = torch.device("cuda:0")
    hbm, src = peaks()
    flush = torch.empty(256 * 1024 * 1024 // 4, device=dev)
    sampler = ClockSampler(0)
    sampler.start()
    rows = []
    for B, cin, cout, P in SHAPES:
        x = torch.randn(B, cin, P, device=dev)
        w = torch.randn(cout, cin, device=dev) / cin ** 0.5
        wb, kp = tcconv.pack_f32(w)
        y = torch.empty(B, cout, P, device=dev)
        bytes_ = 4.0 * B * P * (cin + cout)
        flops = 2.0 * B * P * cin * cout
        kernels = {"tma": lambda: _cabi.call("apb_pwconv_fwd_tf32", [x, wb, None, None, 0, y, B, cin, cout, P, kp], dev)}
        if cout <= 512:
            kernels["rows_r1"] = lambda: _cabi.call("apb_pwconv_fwd_tf32_rows", [x, wb, None, None, 0, y, B, cin, cout, P, kp], dev)
        st = torch.zeros(cout, 2, dtype=torch.float64, device=dev)
        kernels["tma+stats"] = lambda: _cabi.call("apb_pwconv_ex_tf32", [x, wb, None, None, 0, y, 0, st, None, B, cin, cout, P, kp], dev)
        for name, fn in kernels.items():
            us = timeit(fn, flush)
            rows.append({"shape": [B, cin, cout, P], "kernel": name, "us": round(us, 2),
                         "GBps": round(bytes_ / us / 1e3, 1), "hbm_frac": round(bytes_ / us / 1e3 / hbm, 3),
                         "TFLOPs": round(flops / us / 1e6, 1)})
        del x, y
    clocks = sampler.stop()
    for r in rows:
        r["clocks"] = clocks
        r["peak_hbm_GBps"] = hbm
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
