"""Geometry prefetch (geometry.GeometryMailbox, PointNextEncoder.geometry_spec) and the per-forward ball-query memo: the
model computes the same thing whether FPS / ball queries run inline or were prefetched one step ahead on a side stream,
and with the mailbox attached NO sampling or ball-query kernel is launched inside the forward pass."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _data(dev, B, N, seed):
    from adaptpoint_b200.synth import make_clouds
    pos = make_clouds(B, N, "ball", seed).to(dev)
    x = torch.cat([pos, pos[:, :, 1:2] - pos[:, :, 1:2].amin(1, keepdim=True)], -1).transpose(1, 2).contiguous()
    return pos, x


@pytest.mark.parametrize("name,B,N", [("pointnext-b-cls", 2, 4096), ("pointnext-b-seg", 2, 2048)])
def test_prefetched_geometry_equals_inline(cuda_dev, name, B, N):
    from adaptpoint_b200 import _cabi, geometry
    from adaptpoint_b200.models import build_model
    torch.manual_seed(0)
    model = build_model(name).to(cuda_dev).train()
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    pos, x = _data(cuda_dev, B, N, 3)

    def run(mailbox):
        model.zero_grad()
        if mailbox is not None:
            mailbox.attach(pos)
        _cabi.TRACE = []
        out = model({"pos": pos, "x": x})
        torch.cuda.synchronize()
        trace, _cabi.TRACE = _cabi.TRACE, None
        out.float().sum().backward()
        calls = [t[0] for t in trace]
        if hasattr(pos, "_apb_plan"):
            del pos._apb_plan
        g = torch.cat([p.grad.flatten() for p in model.parameters() if p.grad is not None])
        return out.detach().clone(), g.clone(), calls

    want, gwant, calls_inline = run(None)
    geo = torch.cuda.Stream(device=cuda_dev)
    mb = geometry.GeometryMailbox(*model.encoder.geometry_spec(N), geo)
    mb.prefetch(pos)
    torch.cuda.current_stream(cuda_dev).wait_stream(geo)
    got, ggot, calls_pref = run(mb)
    assert calls_inline.count("apb_fps_f32") >= 3 and calls_inline.count("apb_ball_query_f32") >= 4
    # the memo: one ball query per distinct (set, radius) -- fewer than one per block
    nblocks = sum(len(stage) for stage in model.encoder.encoder) - 1
    assert calls_inline.count("apb_ball_query_f32") < nblocks
    assert calls_pref.count("apb_fps_f32") == 0 and calls_pref.count("apb_ball_query_f32") == 0, calls_pref
    # same indices => same arithmetic up to the order of the fp64 statistics atomics (one ulp of mean / rstd)
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-5), (got - want).abs().max()
    assert torch.nn.functional.cosine_similarity(ggot, gwant, dim=0) > 0.9999 and abs(ggot.norm() / gwant.norm() - 1) < 1e-3
    # a second prefetch reuses the static buffers (CUDA-graph replayable)
    ptrs = {k: v.data_ptr() for k, v in mb.nodes.items()}
    mb.prefetch(pos)
    torch.cuda.current_stream(cuda_dev).wait_stream(geo)
    assert ptrs == {k: v.data_ptr() for k, v in mb.nodes.items()}
