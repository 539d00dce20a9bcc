"""Seeded synthetic point clouds: unit-ball-normalised like the reference's
PointCloudCenterAndNormalize (openpoints/transforms/point_transformer_gpu.py:47-61), drawn with a CPU
generator so that every machine sees the same bits."""
import torch


def _normalise(p):
    p = p - p.mean(dim=1, keepdim=True)
    return p / p.norm(dim=-1).amax(dim=1, keepdim=True).unsqueeze(-1).clamp_min(1e-12)


def make_clouds(B, N, kind="ball", seed=0):
    """(B,N,3) float32 CPU tensor.  kinds: ball | shell | quant (1/64 grid: exact distance ties) |
    dup (25% duplicated points) | same (all points identical)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    if kind == "ball":
        d = torch.randn(B, N, 3, generator=g)
        d = d / d.norm(dim=-1, keepdim=True).clamp_min(1e-12)
        r = torch.rand(B, N, 1, generator=g) ** (1.0 / 3.0)
        return _normalise(d * r).contiguous()
    if kind == "shell":
        d = torch.randn(B, N, 3, generator=g)
        d = d / d.norm(dim=-1, keepdim=True).clamp_min(1e-12)
        return _normalise(d + 0.01 * torch.randn(B, N, 3, generator=g)).contiguous()
    if kind == "quant":
        return (torch.randint(-64, 65, (B, N, 3), generator=g).float() / 64.0).contiguous()
    if kind == "dup":
        p = make_clouds(B, N, "ball", seed)
        src = torch.randint(0, N, (B, max(N // 4, 1)), generator=g)
        dst = torch.randint(0, N, (B, max(N // 4, 1)), generator=g)
        for b in range(B):
            p[b, dst[b]] = p[b, src[b]]
        return p.contiguous()
    if kind == "same":
        return torch.full((B, N, 3), 0.25).contiguous()
    raise ValueError(kind)


def scanobjectnn_batch(B=32, N=1024, num_classes=15, seed=0):
    """ScanObjectNN-shaped synthetic batch: pos (B,N,3), x (B,4,N) = xyz + height (y - min y,
    openpoints/dataset/scanobjectnn/scanobjectnn.py:93-96), y (B,) labels."""
    pos = make_clouds(B, N, "shell", seed)
    height = pos[:, :, 1:2] - pos[:, :, 1:2].amin(dim=1, keepdim=True)
    x = torch.cat([pos, height], dim=-1).transpose(1, 2).contiguous()
    g = torch.Generator(device="cpu").manual_seed(seed + 1000)
    y = torch.randint(0, num_classes, (B,), generator=g)
    return {"pos": pos, "x": x, "y": y}
