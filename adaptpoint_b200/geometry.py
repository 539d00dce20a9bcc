"""Sampling plans: the FPS index chains of one batch of clouds, computed ONCE on a side stream.

Farthest-point sampling depends on coordinates only (subsample.py:76-105 takes xyz and nothing else), so every
network that walks the same cloud down the same sample counts gets the same indices: in one AdaptPoint iteration the
imitator (generator_component4_15.py:588-718, reducers 2,2,2,2), the classifier's pass over the real batch
(ganloss_cls.py:31-65) and PointWOLF / the imitator's anchors (fps to 4 points) all sample the SAME loader batch.
A `SamplingPlan` runs those chains once, on its own stream, while the feature work of the consumers proceeds; FPS
launches only 32 CTAs at B=32, so it overlaps with anything.  Consumers call `sample(p, m)`, which returns the
plan's indices when `p` is tagged as a node of a plan holding fps(p, m), and otherwise just runs FPS.

Tags are Python attributes on the tensor objects (`p._apb_plan = (plan, path, version)`): the plan's root is tagged
by its creator, and every `new_p` that `sample` gathers is tagged with the path below it.
"""
import torch

from . import ops


class SamplingPlan:
    def __init__(self, root, chains, stream=None, queries=()):
        """root (B,N,3) f32 contiguous CUDA tensor; chains e.g. [[1024, 512, 256, 128], [4]]: fps(root,1024),
        fps(root[idx],512), ... and fps(root,4).  queries: ball queries between two nodes of the chains,
        (support path, query path, radius, nsample), e.g. ((), (1024,), 0.15, 32).  Work is enqueued on `stream`
        (after what the current stream has enqueued so far), one event per node."""
        self.nodes = {}                       # path tuple -> (idx int64 (B,m), event)
        self.balls = {}                       # (query path, radius, nsample) -> (idx int32 (B,M,K), event)
        self.stream = stream
        cur = torch.cuda.current_stream(root.device)
        base = root.detach()
        if stream is not None:
            stream.wait_stream(cur)
            base.record_stream(stream)
        with torch.cuda.stream(stream if stream is not None else cur):
            coords = {(): base}
            need_xyz = {q[0] for q in queries} | {q[1] for q in queries}
            for chain in chains:
                path = ()
                for k, m in enumerate(chain):
                    parent, path = path, path + (int(m),)
                    if path not in self.nodes:                       # else: shared prefix of two chains
                        idx = ops.furthest_point_sample(coords[parent], int(m)).long()
                        ev = torch.cuda.Event()
                        ev.record()
                        self.nodes[path] = (idx, ev)
                    if (k + 1 < len(chain) or path in need_xyz) and path not in coords:
                        coords[path] = torch.gather(coords[parent], 1, self.nodes[path][0].unsqueeze(-1)
                                                    .expand(-1, -1, 3)).contiguous()
                        for sup, qry, radius, nsample in queries:    # as soon as both ends exist
                            key = (qry, float(radius), int(nsample))
                            if qry == path and sup == parent and key not in self.balls:
                                bidx = ops.ball_query(float(radius), int(nsample), coords[sup], coords[qry])
                                ev = torch.cuda.Event()
                                ev.record()
                                self.balls[key] = (bidx, ev)
        self.adopt(root)

    def adopt(self, tensor, path=()):
        """Declare that `tensor` holds the coordinates of node `path` (the root by default)."""
        tensor._apb_plan = (self, tuple(path), tensor._version)
        return tensor

    def lookup(self, p, m):
        tag = getattr(p, "_apb_plan", None)
        if tag is None or tag[0] is not self or tag[2] != p._version:
            return None
        return self.nodes.get(tag[1] + (int(m),)), tag[1] + (int(m),)


def sample(p, m):
    """(idx int64 (B,m), new_p (B,m,3)) = FPS of p (B,N,3) to m points; new_p is gathered from `p` on the current
    stream (so it stays inside autograd).  Uses the plan `p` is tagged with when it holds this node."""
    tag = getattr(p, "_apb_plan", None)
    idx = None
    if tag is not None and p.is_cuda:
        plan = tag[0]
        hit = plan.lookup(p, m)
        if hit is not None and hit[0] is not None:
            (idx, ev), path = hit
            cur = torch.cuda.current_stream(p.device)
            cur.wait_event(ev)
            idx.record_stream(cur)
    if idx is None:
        idx = ops.furthest_point_sample(p.contiguous(), int(m)).long()
        path = None
    new_p = torch.gather(p, 1, idx.unsqueeze(-1).expand(-1, -1, 3))
    if path is not None:
        tag[0].adopt(new_p, path)
    return idx, new_p


def ball_query(radius, nsample, xyz, new_xyz):
    """ops.ball_query(radius, nsample, xyz, new_xyz), served from the plan when `new_xyz` was sampled from `xyz`
    through `sample` and the plan holds this query."""
    tq, ts = getattr(new_xyz, "_apb_plan", None), getattr(xyz, "_apb_plan", None)
    if (tq is not None and ts is not None and tq[0] is ts[0] and tq[2] == new_xyz._version and ts[2] == xyz._version
            and tq[1][:-1] == ts[1] and len(tq[1]) > 0):
        hit = tq[0].balls.get((tq[1], float(radius), int(nsample)))
        if hit is not None:
            idx, ev = hit
            cur = torch.cuda.current_stream(xyz.device)
            cur.wait_event(ev)
            idx.record_stream(cur)
            return idx
    return ops.ball_query(radius, nsample, xyz, new_xyz)
