"""Op microbench sweep (BASELINE.json configs[2]): FPS / ball_query(r=0.1,K=32) / grouping / gather at
B=32, N=1k..100k, npoint=N/4, plus three_nn / three_interpolate (fwd+bwd) / kNN / deformation+mask (fwd+bwd), ours vs
the reference's CUDA kernels (oracle/_ref) on the same B200, with roofline fractions from the ALGORITHMIC work of
BASELINE.md 2c / SURVEY.md 8d.  SM clocks and throttle reasons are sampled (nvidia-smi, 100 ms) over the whole sweep
and attached to every line.  Writes JSON lines to stdout / --out."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from adaptpoint_b200 import ops  # noqa: E402
from adaptpoint_b200.synth import make_clouds  # noqa: E402
from bench import ClockSampler  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured"
    return 6650.0, "fallback"


def timeit(fn, iters, warm=3, flush=None):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.add_(1.0)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1024,2048,4096,8192,16384,32768,65536,100000")
    ap.add_argument("--B", type=int, default=32)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--ops", default="fps,ball_query,group,group_grad,gather,three_nn,three_interp,knn,deform")
    ap.add_argument("--out", default="")
    ap.add_argument("--tune", action="store_true", help="also sweep the grouping work split (apb_tune_group)")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    hbm, src = peaks()
    refcuda = None
    if not args.no_ref:
        from oracle import refcuda as rc
        refcuda = rc if rc.available() else None
    flush = torch.zeros(160 * 1024 * 1024 // 4, device=dev)  # > 126 MB L2
    B = args.B
    lines = []
    if args.out:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        open(args.out, "w").close()
    sampler = ClockSampler(0)
    sampler.start()
    try:
        sweep(args, dev, hbm, src, refcuda, flush, B, lines)
    finally:
        clocks = sampler.stop()
        print(json.dumps({"clocks": clocks, "note": "nvidia-smi samples over the whole sweep above"}), flush=True)
        if args.out:   # rewrite with the clock record on every line
            with open(args.out, "w") as f:
                for d in lines:
                    d["clocks"] = clocks
                    f.write(json.dumps(d) + "\n")


def sweep(args, dev, hbm, src, refcuda, flush, B, lines):
    for N in [int(s) for s in args.sizes.split(",")]:
        M, K = N // 4, 32
        xyz = make_clouds(B, N, "shell", seed=0).to(dev)
        want = args.ops.split(",")
        idx = ops.furthest_point_sample(xyz, M)
        new_xyz = torch.gather(xyz, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3)).contiguous()
        nbr = ops.ball_query(0.1, K, xyz, new_xyz)

        def emit(op, ms, ms_min, alg_bytes, ref_ms, extra=None):
            d = {"op": op, "B": B, "N": N, "M": M, "K": K, "ms": round(ms, 4), "ms_min": round(ms_min, 4),
                 "alg_GB": round(alg_bytes / 1e9, 4), "achieved_GBs": round(alg_bytes / 1e9 / (ms / 1e3), 1),
                 "frac_hbm": round(alg_bytes / 1e9 / (ms / 1e3) / hbm, 3), "peak": hbm, "peak_src": src,
                 "ref_cuda_ms": None if ref_ms is None else round(ref_ms, 4),
                 "speedup_vs_ref_cuda": None if ref_ms is None else round(ref_ms / ms, 2)}
            if extra:
                d.update(extra)
            lines.append(d)
            print(json.dumps(d), flush=True)
            if args.out:  # append as we go: a later crash keeps what was measured
                with open(args.out, "a") as f:
                    f.write(json.dumps(d) + "\n")

        if "fps" in want:
            ms, mn = timeit(lambda: ops.furthest_point_sample(xyz, M), args.iters, flush=flush)
            rms = None
            if refcuda is not None:
                rms, _ = timeit(lambda: refcuda.fps(xyz, M), max(2, args.iters // 2), warm=1)
            # the 20*N*(m-1) model counts the reference's per-round re-reads, which never leave the SM here: the honest
            # figure of merit is the time per sampling round against its latency floor (one barrier + two CREDUX arg-max
            # + one DSMEM hop for a clustered cloud ~ 0.3 us)
            emit("fps", ms, mn, 20.0 * B * N * (M - 1), rms,
                 {"compulsory_GB": B * (12 * N + 4 * M) / 1e9, "us_per_round": round(ms * 1e3 / max(M - 1, 1), 4),
                  "latency_floor_us_per_round": 0.3})
        if "ball_query" in want:
            ms, mn = timeit(lambda: ops.ball_query(0.1, K, xyz, new_xyz), args.iters, flush=flush)
            rms = None
            if refcuda is not None:
                rms, _ = timeit(lambda: refcuda.ball_query(0.1, K, xyz, new_xyz), max(2, args.iters // 2), warm=1)
            from adaptpoint_b200 import _cabi as _c
            tiled = torch.empty_like(nbr)
            tms, _ = timeit(lambda: _c.call("apb_ball_query_f32", [new_xyz, xyz, tiled, B, N, M, 0.1, K], dev),
                            max(2, args.iters // 2), flush=flush)
            emit("ball_query", ms, mn, 4.0 * B * (3 * N + 3 * M + M * K), rms,
                 {"pairs_upper": B * M * N, "Tpairs_per_s_upper": round(B * M * N / (ms / 1e3) / 1e12, 3),
                  "kernel": "cell grid" if N >= ops.BALL_QUERY_GRID_MIN_N else "tiled O(M*N)",
                  "tiled_kernel_ms": round(tms, 4), "bit_equal_to_tiled": bool(torch.equal(tiled, nbr))})
        for C in (3, 32, 128):
            if C * B * M * K * 4 > 24e9:
                continue
            feat = torch.randn(B, C, N, device=dev)
            # the reference kernels index with 32-bit ints (group_points_gpu.cu:62-71): skip them past 2^31 elements
            ref_ok = refcuda is not None and B * C * M * K < 2 ** 31
            if "group" in want and args.tune:
                from adaptpoint_b200 import _cabi
                for ch, nt in ((1, 256), (1, 512), (1, 1024), (2, 512), (2, 1024), (4, 256), (4, 512), (4, 1024)):
                    if ch > C or ch * N * 4 > 200 * 1024:
                        continue
                    _cabi.tune_group(ch, nt)
                    ms, mn = timeit(lambda: ops.grouping_operation(feat, nbr), args.iters, flush=flush)
                    emit(f"group_C{C}@ch{ch}nt{nt}", ms, mn, 4.0 * B * (C * N + M * K + C * M * K), None)
                    if C in (3, 128):
                        ms, mn = timeit(lambda: ops.gather_operation(feat, idx), args.iters, flush=flush)
                        emit(f"gather_C{C}@ch{ch}nt{nt}", ms, mn, 4.0 * B * (2 * C * M + M), None)
                _cabi.tune_group(0, 0)
            if "group" in want:
                ms, mn = timeit(lambda: ops.grouping_operation(feat, nbr), args.iters, flush=flush)
                rms = None
                if ref_ok:
                    rms, _ = timeit(lambda: refcuda.group(feat, nbr), max(2, args.iters // 2), warm=1)
                extra = None
                if N * 4 > 200 * 1024:             # cluster (DSMEM) kernel vs the direct L2 gather it replaces
                    from adaptpoint_b200 import _cabi
                    _cabi.tune_group(-1, 0)
                    ms_d, _ = timeit(lambda: ops.grouping_operation(feat, nbr), max(2, args.iters // 2), flush=flush)
                    _cabi.tune_group(0, 0)
                    extra = {"ms_direct_l2_gather": round(ms_d, 4)}
                emit(f"group_C{C}", ms, mn, 4.0 * B * (C * N + M * K + C * M * K), rms, extra)
            if "group" in want and C >= 32:     # bf16 features: half the feature bytes, 16-byte stores of eight values
                fb = feat.to(torch.bfloat16)
                ms, mn = timeit(lambda: ops.grouping_operation_bf16(fb, nbr), args.iters, flush=flush)
                emit(f"group_bf16_C{C}", ms, mn, 2.0 * B * (C * N + C * M * K) + 4.0 * B * M * K, None)
                del fb
            if "group_grad" in want:
                go = torch.randn(B, C, M, K, device=dev)
                gp = torch.zeros(B, C, N, device=dev)
                from adaptpoint_b200 import _cabi

                def bwd():
                    gp.zero_()
                    _cabi.call("apb_group_grad_f32", [go, nbr, gp, B, C, N, M, K], dev)
                ms, mn = timeit(bwd, args.iters, flush=flush)
                rms = None
                if ref_ok:
                    rms, _ = timeit(lambda: refcuda.group_grad(go, nbr, N), max(2, args.iters // 2), warm=1)
                _cabi.tune_group_grad(1)           # the reference's scheme in our code: scattered L2 reductions
                ms1, _ = timeit(bwd, max(2, args.iters // 2), flush=flush)
                _cabi.tune_group_grad(0)
                emit(f"group_grad_C{C}", ms, mn, 4.0 * B * (C * M * K + M * K + C * N), rms,
                     {"ms_l2_reductions": round(ms1, 4)})
                del go, gp
            if "gather" in want and C in (3, 128):
                ms, mn = timeit(lambda: ops.gather_operation(feat, idx), args.iters, flush=flush)
                rms = None
                if refcuda is not None:
                    rms, _ = timeit(lambda: refcuda.gather(feat, idx), max(2, args.iters // 2), warm=1)
                emit(f"gather_C{C}", ms, mn, 4.0 * B * (2 * C * M + M), rms)
            del feat
        # ---- decoder-side ops (S3DIS-shaped levels: n unknown points interpolate from m = n/4 known ones) ----------
        from adaptpoint_b200 import _cabi
        n_unk, m_kn = N, M
        if "three_nn" in want:
            ms, mn = timeit(lambda: ops.three_nn(xyz, new_xyz), args.iters, flush=flush)
            rms = None
            if refcuda is not None:
                rms, _ = timeit(lambda: refcuda.three_nn(xyz, new_xyz), max(2, args.iters // 2), warm=1)
            emit("three_nn", ms, mn, 4.0 * B * (3 * n_unk + 3 * m_kn + 6 * n_unk), rms,
                 {"pairs": B * n_unk * m_kn, "Tpairs_per_s": round(B * n_unk * m_kn / (ms / 1e3) / 1e12, 3),
                  "bound": "fp32 ALU pair rate (~5.3 T pairs/s), bytes negligible"})
        if "three_interp" in want:
            dist, tidx = ops.three_nn(xyz, new_xyz)
            wgt = 1.0 / (dist + 1e-8)
            wgt = (wgt / wgt.sum(dim=2, keepdim=True)).contiguous()
            for C in (32, 128):
                if C * B * n_unk * 4 > 24e9:
                    continue
                kf = torch.randn(B, C, m_kn, device=dev)
                ms, mn = timeit(lambda: ops.three_interpolate(kf, tidx, wgt), args.iters, flush=flush)
                rms = None
                if refcuda is not None and B * C * n_unk < 2 ** 31:
                    rms, _ = timeit(lambda: refcuda.three_interp(kf, tidx, wgt), max(2, args.iters // 2), warm=1)
                emit(f"three_interp_C{C}", ms, mn, 4.0 * B * (C * m_kn + 6 * n_unk + C * n_unk), rms)
                go = torch.randn(B, C, n_unk, device=dev)
                gp = torch.zeros(B, C, m_kn, device=dev)

                def ibwd():
                    gp.zero_()
                    _cabi.call("apb_three_interp_grad_f32", [go, tidx, wgt, gp, B, C, n_unk, m_kn], dev)
                ms, mn = timeit(ibwd, args.iters, flush=flush)
                rms = None
                if refcuda is not None and B * C * n_unk < 2 ** 31:
                    rms, _ = timeit(lambda: refcuda.three_interp_grad(go, tidx, wgt, m_kn), max(2, args.iters // 2), warm=1)
                emit(f"three_interp_grad_C{C}", ms, mn, 4.0 * B * (C * m_kn + 6 * n_unk + C * n_unk), rms)
                del kf, go, gp
        if "knn" in want and N <= 32768:
            ms, mn = timeit(lambda: ops.knn_query(K, xyz, new_xyz), args.iters, flush=flush)
            tms = None
            if N <= 16384:      # the (B,N,M) distance matrix of the reference's formulation: 8.6 GB at N = 16384
                tms, _ = timeit(lambda: torch.cdist(xyz, new_xyz).topk(K, dim=1, largest=False), max(2, args.iters // 2), warm=1)
            emit("knn", ms, mn, 4.0 * B * (3 * N + 3 * M + 2 * M * K), None,
                 {"pairs": B * M * N, "Tpairs_per_s": round(B * M * N / (ms / 1e3) / 1e12, 3),
                  "torch_cdist_topk_ms": None if tms is None else round(tms, 4),
                  "speedup_vs_torch_cdist_topk": None if tms is None else round(tms / ms, 2),
                  "bound": "fp32 ALU pair rate; the reference's kNN is torch.cdist + topk (group.py:12-28)"})
        if "deform" in want:
            from adaptpoint_b200.imitator import deform_mask
            A = 4
            anchor = new_xyz[:, :A].contiguous()
            aff = torch.randn(B, A, 12, device=dev) * 0.1
            aff[:, :, 0] += 1; aff[:, :, 4] += 1; aff[:, :, 8] += 1
            aff.requires_grad_(True)
            proj = torch.ones(B, 3, device=dev)
            mask = (torch.rand(B, N, device=dev) > 0.1).float().requires_grad_(True)
            ms, mn = timeit(lambda: deform_mask(xyz, anchor, aff, proj, mask, 0.5), args.iters, flush=flush)
            emit("deform_mask_fwd", ms, mn, 4.0 * B * (3 * N + 2 * N + 3 * N), None)
            out = deform_mask(xyz, anchor, aff, proj, mask, 0.5)
            go = torch.randn_like(out)
            ms, mn = timeit(lambda: torch.autograd.grad(out, (aff, mask), go, retain_graph=True), args.iters, flush=flush)
            emit("deform_mask_bwd", ms, mn, 4.0 * B * (3 * N + 2 * N + 3 * N + N), None)


if __name__ == "__main__":
    main()
