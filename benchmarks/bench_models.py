#!/usr/bin/env python
"""Data-parallel training-step throughput of the wider PointNeXt configurations (BASELINE.json configs[3], configs[4]).

  python benchmarks/bench_models.py [--gpus N] [--steps K] [--warmup W] [--cases b-cls,l-cls,xl-seg]
  (N>1: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 ... benchmarks/bench_models.py --gpus N)

One step = forward + loss + backward + flat NCCL gradient all-reduce (train.allreduce_grads) + AdamW update of one
per-GPU batch held in HBM.  By default the step is captured once as ONE CUDA graph (side-stream warm-up, capturable
fused AdamW, the NCCL all-reduce inside the capture: the pattern of train.AdaptPointTrainer.capture_graph) and
replayed; `--eager` launches it kernel by kernel (the plain `build_model(...)` path, host-launch bound: round 1's
0.85 / 0.63 at N=8).  Per-GPU batches are SURVEY.md 8d's: 32 clouds of 8192 points (B, L), 8 clouds of 24000 points
(XL); `--batch` overrides.  The gradient all-reduce is one flat collective after the backward pass: 15-166 MB over
NVLink is 0.1-0.5 ms of a 30-120 ms step, so bucketing it under the backward pass would buy < 0.5 %.
Timing: CUDA events around the K timed steps, barrier + synchronize on both sides, max over ranks; samples/s is the
whole job (weak scaling, per-GPU batch fixed); SM clocks / throttle reasons sampled (nvidia-smi, 100 ms) during the
timed region of every case.  One JSON line per case on rank 0.
Inputs are synthetic (seeded unit-ball clouds, xyz + height), weights random-init.  bench.py stays the driver's
contract (configs[1]); this file only records configs[3]/[4] beside it.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

# name -> (CONFIGS key, per-GPU batch, points, classes); shapes of BASELINE.json configs[3] / configs[4]
CASES = {
    "b-cls": ("pointnext-b-cls", 32, 8192, 40),
    "l-cls": ("pointnext-l-cls", 32, 8192, 40),
    "xl-seg": ("pointnext-xl-seg", 8, 24000, 13),
}


def make_batch(B, N, classes, seg, seed, dev):
    from adaptpoint_b200.synth import make_clouds
    pos = make_clouds(B, N, "ball", seed)
    h = pos[:, :, 1:2] - pos[:, :, 1:2].amin(dim=1, keepdim=True)
    x = torch.cat([pos, h], -1).transpose(1, 2).contiguous()
    g = torch.Generator().manual_seed(seed + 77)
    y = torch.randint(0, classes, (B, N) if seg else (B,), generator=g)
    return pos.to(dev), x.to(dev), y.to(dev)


def run_case(name, args, rank, world, dev):
    from adaptpoint_b200 import tcconv
    from adaptpoint_b200.models import build_model
    from adaptpoint_b200.train import allreduce_grads
    key, B, N, classes = CASES[name]
    if args.batch:
        B = args.batch
    seg = key.endswith("seg")
    torch.manual_seed(0)                       # same initial weights on every rank
    model = build_model(key).to(dev).train()
    params = [p for p in model.parameters() if p.requires_grad]
    opt = torch.optim.AdamW(params, lr=1e-3, weight_decay=1e-4, fused=True, capturable=args.graph)
    pos, x, y = make_batch(B, N, classes, seg, 100 + rank, dev)

    # Geometry prefetch (geometry.GeometryMailbox): FPS is strictly sequential (6000 rounds for 24000 -> 6000) and needs
    # coordinates only, so the FPS / ball-query chain of the NEXT batch runs on a side stream beside the backward pass of
    # the current one, as a loader-driven training loop would arrange it.  Every step still executes one full chain
    # inside the timed region (for the batch after it); --no-prefetch keeps it inline in the forward pass.
    mailbox = None
    if args.prefetch:
        from adaptpoint_b200 import geometry
        geo = torch.cuda.Stream(device=dev)
        mailbox = geometry.GeometryMailbox(*model.encoder.geometry_spec(N), geo)
        mailbox.prefetch(pos)
        torch.cuda.current_stream(dev).wait_stream(geo)

    def step():
        opt.zero_grad(set_to_none=True)
        with tcconv.conv_backend(args.conv):   # 'tc': 1x1 convs (fwd, dgrad, wgrad) on libapb200's tcgen05 kernels
            if mailbox is not None:
                mailbox.attach(pos)
            if seg:
                loss = torch.nn.functional.cross_entropy(model({"pos": pos, "x": x}), y)
            else:
                _, loss = model.get_logits_loss({"pos": pos, "x": x}, y)
            if mailbox is not None:
                mailbox.prefetch(pos)          # the next batch (synthetic: the same clouds) -- beside the backward pass
            loss.backward()
        allreduce_grads(params)
        opt.step()
        if mailbox is not None:
            torch.cuda.current_stream(dev).wait_stream(mailbox.stream)
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.dump_kernels and rank == 0:       # per C-ABI call CUDA-event trace of two eager steps (host-launch skew included)
        from adaptpoint_b200 import _cabi
        for _ in range(2):
            step()
        torch.cuda.synchronize()
        _cabi.TRACE = []
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(2):
            step()
        e1.record()
        torch.cuda.synchronize()
        trace, _cabi.TRACE = _cabi.TRACE, None
        agg = {}
        for nm, a, a0, a1 in trace:
            d = agg.setdefault(nm, [0, 0.0])
            d[0] += 1
            d[1] += a0.elapsed_time(a1)
        rows = sorted(({"call": k, "launches_per_step": v[0] // 2, "ms_per_step": round(v[1] / 2, 3)} for k, v in agg.items()),
                      key=lambda r: -r["ms_per_step"])
        with open(f"{args.dump_kernels}.{name}.json", "w") as f:
            json.dump({"eager_ms_per_step": round(e0.elapsed_time(e1) / 2, 3), "ours_ms_per_step": round(sum(r["ms_per_step"] for r in rows), 3),
                       "calls": rows}, f, indent=1)
    graph = None
    if args.graph:
        eager_step = step
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3):                 # real steps: allocator, cuDNN/NCCL and optimizer state settle
                eager_step()
        torch.cuda.current_stream(dev).wait_stream(side)
        barrier()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            g_loss = eager_step()

        def step():
            graph.replay()
            return g_loss

    from bench import ClockSampler
    # nvidia-smi needs ~0.1 s to deliver its first row and a timed region can be shorter than that (10 steps of 6 ms): the
    # sampler runs from the first warm-up replay on (same graph, same load) through the timed steps
    sampler = ClockSampler(dev.index)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        loss = step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None:
        clocks["window"] = "warm-up + timed replays"
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step = ms.item() / args.steps
    peak_gb = torch.cuda.max_memory_allocated() / 2**30
    if rank == 0:
        print(json.dumps({
            "metric": f"pointnext_{name.replace('-', '_')}_train_samples_per_s", "value": round(B * world / ms_step * 1e3, 2),
            "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "clocks": clocks,
            "ms_per_step": round(ms_step, 3), "higher_is_better": True, "scaling": "weak",
            "dtype": "tf32" if args.conv == "tc" else "f32",
            "data": "synthetic", "loss": round(float(loss.detach()), 4), "hbm_peak_alloc_gib": round(peak_gb, 2),
            "params_m": round(sum(p.numel() for p in params) / 1e6, 3),
            "config": {"workload": f"{key} fwd+bwd+allreduce+AdamW, B={B}/GPU, N={N}, {classes} classes, "
                                   f"{'one CUDA graph' if args.graph else 'eager'}, "
                                   f"geometry {'of the next batch prefetched beside the backward pass' if args.prefetch else 'inline'}, "
                                   f"1x1 convs: {'libapb200 tcgen05 tf32' if args.conv == 'tc' else 'cuDNN fp32'}",
                       "parallelism": f"dp{world}"}}), flush=True)
    if graph is not None:                      # a graph holding NCCL work must go before destroy_process_group
        torch.cuda.synchronize()
        graph.reset()
        graph = None
    del model, opt
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cases", default="b-cls,l-cls,xl-seg")
    ap.add_argument("--conv", default="tc", choices=["tc", "torch"])
    ap.add_argument("--eager", dest="graph", action="store_false", help="launch the step kernel by kernel instead of "
                    "replaying one CUDA graph")
    ap.add_argument("--graph", dest="graph", action="store_true", help="(default) capture the step in one CUDA graph")
    ap.add_argument("--dump-kernels", default="", help="prefix of per-case JSON files with a per-C-ABI-call time table")
    ap.add_argument("--no-prefetch", dest="prefetch", action="store_false", help="FPS / ball queries inline in the "
                    "forward pass instead of one step ahead on a side stream")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch override (default: SURVEY 8d's 32 / 32 / 8)")
    ap.set_defaults(graph=True)
    args = ap.parse_args()
    if not torch.cuda.is_available():
        raise SystemExit("bench_models.py needs a CUDA device (no CPU fallback)")
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    try:
        for name in args.cases.split(","):
            run_case(name, args, rank, world, dev)
    finally:
        if world > 1:
            torch.cuda.synchronize()
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
