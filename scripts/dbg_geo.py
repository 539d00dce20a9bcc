import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from adaptpoint_b200 import geometry, ops
from adaptpoint_b200.models import build_model
from adaptpoint_b200.synth import make_clouds
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = build_model("pointnext-b-cls").to(dev).train()
N = 4096
pos = make_clouds(2, N, "ball", 3).to(dev)
x = torch.cat([pos, pos[:, :, 1:2] - pos[:, :, 1:2].amin(1, keepdim=True)], -1).transpose(1, 2).contiguous()
chains, queries = model.encoder.geometry_spec(N)
print("chains", chains); print("queries", queries)
geo = torch.cuda.Stream(device=dev)
mb = geometry.GeometryMailbox(chains, queries, geo)
mb.prefetch(pos); torch.cuda.current_stream().wait_stream(geo); torch.cuda.synchronize()
print("mailbox nodes", list(mb.nodes), "balls", list(mb.balls))
orig_bq, orig_sample = geometry.ball_query, geometry.sample
log = []
def bq(radius, nsample, xyz, new_xyz):
    out = orig_bq(radius, nsample, xyz, new_xyz)
    want = ops.ball_query(radius, nsample, xyz.contiguous(), new_xyz.contiguous())
    log.append(("bq", radius, nsample, tuple(xyz.shape), tuple(new_xyz.shape), bool(torch.equal(out, want)), getattr(xyz, "_apb_plan", (None, None))[1], getattr(new_xyz, "_apb_plan", (None, None))[1]))
    return out
def sm(p, m):
    idx, new_p = orig_sample(p, m)
    want = ops.furthest_point_sample(p.contiguous(), m).long()
    log.append(("fps", m, bool(torch.equal(idx, want)), getattr(p, "_apb_plan", (None, None))[1]))
    return idx, new_p
import adaptpoint_b200.layers as L, adaptpoint_b200.pointnext as P, adaptpoint_b200.fused_train as FT
geometry.ball_query = bq; geometry.sample = sm
mb.attach(pos)
out = model({"pos": pos, "x": x})
torch.cuda.synchronize()
for l in log: print(l)
